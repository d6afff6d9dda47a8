import sys, numpy as np, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import common
from common import A
names=["dL_dmeans2D","dL_dopacity","dL_dmeans3D","dL_dcov3D","dL_dsh","dL_dscales","dL_drotations","dL_dcolors"]
def stat(ga,gb):
    out=[]
    for k in names:
        if k not in ga or k not in gb: continue
        a=A(ga[k]).astype(np.float64).reshape(-1); b=A(gb[k]).astype(np.float64).reshape(-1)
        if a.size==0: continue
        mx=max(np.abs(a).max(),np.abs(b).max())+1e-30
        d=np.abs(a-b)
        # smallest floor factor f such that d <= 1e-3*max(|a|,|b|) + f*mx
        f=((d-1e-3*np.maximum(np.abs(a),np.abs(b)))/mx).max()
        out.append(f"{k[3:]}:{f:.1e}")
    return " ".join(out)
for case in ["tiny","small","ragged","inside","dense","deg1","precomp","obj"]:
    sc,cam=common.make_case(case); dp=common.upstream_gradient(cam)
    o=common.run_ours(sc,cam,dp); r=common.run_reference(sc,cam,dp); c=common.run_oracle(sc,cam,dp)
    print(case,"\n  ours-orc",stat(o["grads"],c["grads"]),"\n  ref-orc ",stat(r["grads"],c["grads"]),"\n  ours-ref",stat(o["grads"],r["grads"]))
    img_o=A(o["out_color"]); img_c=c["out_color"]; d=np.abs(img_o-img_c)
    print("  img ours-orc max",d.max(),"n>1e-4:",(d>1e-4).sum(), " ref-orc max", np.abs(A(r["out_color"])-img_c).max())
