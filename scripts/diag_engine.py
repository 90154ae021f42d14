import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from dpt_b200.models.resnet import build_model
def cos(a,b):
    a,b=a.flatten().float(),b.flatten().float(); return float(torch.dot(a,b)/(a.norm()*b.norm()+1e-12))
torch.manual_seed(0)
dev=torch.device("cuda")
import sys as _s; ARCH=_s.argv[1] if len(_s.argv)>1 else "resnet18"; IMG=int(_s.argv[2]) if len(_s.argv)>2 else 32; BS=int(_s.argv[3]) if len(_s.argv)>3 else 32
model=build_model(dev,ARCH,10)
if ARCH=="resnet50":
    with torch.no_grad():
        for n,p in model.named_parameters():
            if n.endswith("bn3.weight"): p.fill_(0.1); ref=copy.deepcopy(model); ref.native="never"; model.native="always"
ref2=copy.deepcopy(ref)
x=torch.randn(BS,3,IMG,IMG,device=dev); t=torch.randint(0,10,(BS,),device=dev)
out=model(x); F.cross_entropy(out,t).backward()
out_r=ref(x); F.cross_entropy(out_r,t).backward()
with torch.autocast("cuda",dtype=torch.bfloat16):
    out2=ref2(x); l2=F.cross_entropy(out2,t)
l2.backward()
for (n,p),q,r in zip(model.named_parameters(),ref.parameters(),ref2.parameters()):
    flag = "  <<<<<<" if cos(p.grad,q.grad) < cos(r.grad,q.grad)-0.05 else ""
    print(f"{n:40s}{flag} ours-vs-fp32 {cos(p.grad,q.grad):.4f}   torch-bf16-vs-fp32 {cos(r.grad,q.grad):.4f}  norm ours {p.grad.norm():.4g} ref {q.grad.norm():.4g}")
