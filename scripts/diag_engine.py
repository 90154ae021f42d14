import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from dpt_b200.models.resnet import build_model
def cos(a,b):
    a,b=a.flatten().float(),b.flatten().float(); return float(torch.dot(a,b)/(a.norm()*b.norm()+1e-12))
torch.manual_seed(0)
dev=torch.device("cuda")
model=build_model(dev,"resnet18",10); ref=copy.deepcopy(model); ref.native="never"; model.native="always"
ref2=copy.deepcopy(ref)
x=torch.randn(32,3,32,32,device=dev); t=torch.randint(0,10,(32,),device=dev)
out=model(x); F.cross_entropy(out,t).backward()
out_r=ref(x); F.cross_entropy(out_r,t).backward()
with torch.autocast("cuda",dtype=torch.bfloat16):
    out2=ref2(x); l2=F.cross_entropy(out2,t)
l2.backward()
for (n,p),q,r in zip(model.named_parameters(),ref.parameters(),ref2.parameters()):
    print(f"{n:40s} ours-vs-fp32 {cos(p.grad,q.grad):.4f}   torch-bf16-vs-fp32 {cos(r.grad,q.grad):.4f}  norm ours {p.grad.norm():.4g} ref {q.grad.norm():.4g}")
