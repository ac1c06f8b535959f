import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from conftest import load_pkg
from oracle import bn254 as bn, binfile
pkg=load_pkg(); G=bn.G1
tab={G.mul(bn.G1_GEN,k):k for k in range(1,2000)}
def chk(ks, op):
    pts=[G.mul(bn.G1_GEN,k) if k else None for k in ks]
    raw=b"".join(binfile.enc_g1(p) for p in pts)
    out=pkg.curve_op(False,op,raw,len(pts))
    got=binfile.dec_g1(out)
    print(op, ks[:6], tab.get(got, got), 'oncurve', got is None or G.is_on_curve(got))
for ks in ([5],[5,7],[5,5],[1,2,3],[9]*128):
    chk(ks,1)
