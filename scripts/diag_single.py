import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from dpt_b200.models.resnet import resnet18
from dpt_b200.optim import SGD
dev=torch.device("cuda")
torch.manual_seed(100)
base=resnet18(10).to(dev); base.native="never"
ours=copy.deepcopy(base); ref=copy.deepcopy(base)
oo=SGD(ours.parameters(),lr=0.002,momentum=0.9,weight_decay=5e-4)
o1=torch.optim.SGD(ref.parameters(),lr=0.002,momentum=0.9,weight_decay=5e-4)
names=[n for n,_ in ours.named_parameters()]
def rel(a,b): return ((a-b).abs().max()/(b.abs().max()+1e-30)).item()
g=torch.Generator(device=dev).manual_seed(7)
for it in range(4):
    x=torch.randn(16,3,32,32,device=dev,generator=g); t=torch.randint(0,10,(16,),device=dev,generator=g)
    for m,o in ((ours,oo),(ref,o1)):
        o.zero_grad(set_to_none=True); F.cross_entropy(m(x),t).backward()
    torch.cuda.synchronize()
    print(it, "grad relerr worst", max((rel(a.grad,b.grad),n) for n,a,b in zip(names,ours.parameters(),ref.parameters())))
    print(it, "buffers relerr worst", max((rel(a.float(),b.float()),n) for (n,a),b in zip(ours.named_buffers(),ref.buffers())))
    oo.step(); o1.step(); torch.cuda.synchronize()
    print(it, "param relerr worst", max((rel(a,b),n) for n,a,b in zip(names,ours.parameters(),ref.parameters())))
