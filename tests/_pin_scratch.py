import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, time
from intercession_b200 import scenes, abi
import refs
def fields_equal(a,b,prefix=""):
    bad=[]
    if a.shape!=b.shape: return [("shape",a.shape,b.shape)]
    for name in a.dtype.names:
        x,y=a[name],b[name]
        if x.dtype.names: bad+=fields_equal(x,y,prefix+name+".")
        else:
            xe = x.view(np.uint32) if x.dtype==np.float32 else x
            ye = y.view(np.uint32) if y.dtype==np.float32 else y
            if not np.array_equal(xe,ye):
                idx=np.argwhere(xe!=ye)
                bad.append((prefix+name, len(idx), idx[:3].tolist(), x[tuple(idx[0])], y[tuple(idx[0])]))
    return bad
def cmp(a,b):
    out=[]
    for k in ("T","P","C","hits","ints"):
        out+=fields_equal(a[k],b[k])
    for k in ("tstate","tcoh"):
        if not np.array_equal(a[k],b[k]): out.append(k)
    return out
if __name__=="__main__":
    for seed in range(3):
        a = scenes.arena(300, seed=seed, fill_side=9.0, spacing=0.9, jitter=0.2, kinds=("sphere","box","aabb"), temporal=True)
        a.ids[10:160] |= 1; a.ids[10:160] &= ~np.uint32(14)
        a.tstate[100:200]=abi.TS_FORKED
        for host in (0,14):
          for flags in (0,1):
            r = refs.procedures_tick(refs.ref(), "icxref", a, host=host, coherency=7, flags=flags)
            o = refs.procedures_tick(refs.oracle(), "icxo", a, host=host, coherency=7, flags=flags)
            print(seed, host, flags, len(r.hits), len(r.ints), cmp(r,o))
    s = scenes.case_study(push=True)
    r=s; o=s
    for t in range(300):
        r = refs.procedures_tick(refs.ref(), "icxref", r, coherency=t)
        o = refs.procedures_tick(refs.oracle(), "icxo", o, coherency=t)
    print("case", cmp(r,o), len(r.hits), r.T["origin"][2:])
