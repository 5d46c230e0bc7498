import sys, numpy as np, ctypes
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import oraclelib
chk = oraclelib.oracle(); chk.lib.fsqo_set_exact_rmq.argtypes=[ctypes.c_int]
rng=np.random.default_rng(7)
best=None
for trial in range(3000):
    m=int(rng.integers(4,9)); n=int(rng.integers(700,1600)); L=int(rng.integers(2,6)); F=int(rng.integers(2,4))
    founders=rng.integers(0,2,size=(F,n)).astype(np.uint8)
    rows=np.empty((m,n),np.uint8)
    for r in range(m):
        f=int(rng.integers(0,F)); pos=0
        while pos<n:
            ln=int(rng.geometric(1/ rng.choice([150,400,900]))); rows[r,pos:pos+ln]=founders[f,pos:pos+ln]; pos+=ln; f=int(rng.integers(0,F))
    mut=rng.random((m,n))< rng.choice([0.0, 0.002, 0.01])
    rows=np.ascontiguousarray(np.frombuffer(b"AC",np.uint8)[np.where(mut,1-rows,rows)])
    res=[]
    for ex in (0,1):
        chk.lib.fsqo_set_exact_rmq(ex)
        with chk.open(rows) as o:
            st=o.segment(L); res.append((st,[x.copy() for x in o.dp()],[x.copy() for x in o.traceback()],[x.copy() for x in o.reduced()] if st==0 else None))
    chk.lib.fsqo_set_exact_rmq(0)
    if res[0][0]!=0: continue
    tb_diff = len(res[0][2][0])!=len(res[1][2][0]) or (res[0][2][0]!=res[1][2][0]).any()
    red_diff = len(res[0][3][0])!=len(res[1][3][0]) or (res[0][3][1]!=res[1][3][1]).any()
    mx=(res[0][1][2]!=res[1][1][2]).sum()
    if red_diff:
        score=(mx>0, n)
        print(trial,m,n,L,"tb",tb_diff,"red",red_diff,"maxdiffs",mx, "segs", len(res[0][3][0]), len(res[1][3][0]))
        if best is None or score>best[0]:
            best=(score,rows.copy(),L)
        if mx>0: break
if best: np.save('/tmp/quirk_rows.npy',best[1]); open('/tmp/quirk_L','w').write(str(best[2])); print("saved",best[1].shape,best[2])
