import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from conftest import load_pkg
from oracle import bn254 as bn, binfile
pkg=load_pkg(); G=bn.G1
def chk(ks, op):
    pts=[G.mul(bn.G1_GEN,k) if k else None for k in ks]
    raw=b"".join(binfile.enc_g1(p) for p in pts)
    got=binfile.dec_g1(pkg.curve_op(False,op,raw,len(pts)))
    want=G.mul(bn.G1_GEN,sum(ks)%bn.R)
    print(op, ks[:6], 'OK' if got==want else 'FAIL')
for ks in ([5],[5,7],[5,5],[5,7,11],[5,7,11,13],[1,2,3,5,2],[5,0,7],[3]*4,[9]*128,[9]*129, list(range(1,200))):
    chk(ks,1); chk(ks,0)
pts=[G.mul(bn.G1_GEN,k) for k in (1,2,3,77)]
out=pkg.curve_op(False,2,b"".join(binfile.enc_g1(p) for p in pts),4)
print([binfile.dec_g1(out[64*i:64*i+64])==G.mul(p,2) for i,p in enumerate(pts)])
