"""Is the RPP result of corner ordering 1 the exact mirror (R * diag(1,-1,-1), same t) of ordering 0's?"""
import ctypes, os, sys
from multiprocessing import Pool
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT+'/tests'); sys.path.insert(0, ROOT+'/tools/rpp_emul')
from frames import K_for, poses
from oracle import pyoracle as po
import compare as cmp
def main():
    n=int(sys.argv[1]) if len(sys.argv)>1 else 64
    emul=cmp.build()
    dp=lambda a: a.ctypes.data_as(ctypes.c_void_p)
    with Pool(os.cpu_count()) as pool:
        allq=pool.map(cmp.quads_of, [(i,p,640,480) for i,p in enumerate(poses(n, 0x4D414E54+2))])
    s=0.16
    poss=[[(s,s),(-s,s),(-s,-s),(s,-s)],[(s,-s),(-s,-s),(-s,s),(s,s)]]
    D=np.diag([1.0,-1.0,-1.0])
    nq=0; dev_exact=0; dev_bad=0; dev_okdiff=0; orc_max=0.0; orc_itdiff=0; orc_okdiff=0; dev_its_diff=0
    olib=po.lib()
    dbg_it=(ctypes.c_int*8).in_dll(olib,"mo_rpp_debug_iters"); dbg_n=ctypes.c_int.in_dll(olib,"mo_rpp_debug_n")
    for fi,qs in allq:
        for qi,q in enumerate(qs):
            rays=np.ascontiguousarray(q,np.float64)
            out=[]
            oro=[]
            for e in range(2):
                R9,t3,its=np.zeros(9),np.zeros(3),(ctypes.c_int*8)()
                nit,eo,ei=ctypes.c_int(),ctypes.c_double(),ctypes.c_double()
                ok=emul.emul_rpp(dp(rays),ctypes.c_double(s),e,dp(R9),dp(t3),its,ctypes.byref(nit),ctypes.byref(eo),ctypes.byref(ei))
                out.append((ok,R9.reshape(3,3).copy(),t3.copy(),[its[k] for k in range(nit.value)],eo.value,ei.value))
                model=np.zeros((3,4)); ip=np.ones((3,4))
                for k in range(4):
                    model[0,k],model[1,k]=poss[e][k]; ip[0,k],ip[1,k]=q[2*k],q[2*k+1]
                oro.append(po.rpp(model,ip)+([dbg_it[k] for k in range(dbg_n.value)],))
            nq+=1
            (ok0,R0,t0,it0,oe0,ie0),(ok1,R1,t1,it1,oe1,ie1)=out
            if bool(ok0)!=bool(ok1): dev_okdiff+=1; continue
            if ok0:
                if np.array_equal(R1,R0@D) and np.array_equal(t1,t0) and oe0==oe1 and ie0==ie1: dev_exact+=1
                else:
                    dev_bad+=1
                    if dev_bad<=5: print("device differs", fi, qi, np.abs(R1-R0@D).max(), np.abs(t1-t0).max(), it0, it1)
                if sorted(it0[1:])!=sorted(it1[1:]) or it0[0]!=it1[0]: dev_its_diff+=1
            o0,o1=oro
            if o0[0]!=o1[0]: orc_okdiff+=1
            elif o0[0]:
                orc_max=max(orc_max, np.abs(o1[1]-o0[1]@D).max(), np.abs(o1[2]-o0[2]).max())
                if sorted(o0[-1][1:])!=sorted(o1[-1][1:]) or o0[-1][0]!=o1[-1][0]: orc_itdiff+=1
    print(f"{n} frames, {nq} quads: device ordering-1 result == mirror of ordering 0 bit for bit in {dev_exact}, differs in {dev_bad}, ok flag differs {dev_okdiff}, iteration counts differ {dev_its_diff}")
    print(f"oracle: max |ordering 1 - mirror(ordering 0)| = {orc_max:.3e}, iteration counts differ in {orc_itdiff}, ok differs {orc_okdiff}")
main()
