import sys, torch
sys.path.insert(0, "/root/repo")
from stereoformer_b200 import ops
torch.manual_seed(0)
for shape in [(8,256,68,120,24),(2,64,37,200,48),(1,32,5,40,8),(4,96,33,260,40),(3,256,50,64,24)]:
    B,C,H,W,D=shape
    L=torch.randn(B,C,H,W,device="cuda"); R=torch.randn(B,C,H,W,device="cuda")
    ref=torch.zeros(B,D,H,W,device="cuda",dtype=torch.float64)
    for d in range(D):
        ref[:,d,:,d:]=(L.double()[:,:,:,d:]*R.double()[:,:,:,:W-d]).mean(1)
    o0=ops.corr_volume(L,R,D)
    bad=0
    for i in range(200):
        o=ops.corr_volume(L,R,D)
        bad+=int(not torch.equal(o,o0))
    print(shape,"max err",(o0.double()-ref).abs().max().item(),"nondeterministic runs",bad)
