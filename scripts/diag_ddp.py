import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist, torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as TorchDDP
rank=int(os.environ["RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev=torch.device("cuda",rank)
dist.init_process_group("nccl", device_id=dev)
from dpt_b200.models.resnet import resnet18
from dpt_b200.optim import SGD
from dpt_b200.parallel.ddp import DistributedDataParallel as OurDDP
torch.manual_seed(100+rank)
base=resnet18(10).to(dev); base.native="never"
ours=OurDDP(copy.deepcopy(base),device_ids=[rank],backend=os.environ.get("DIAG_BACKEND","native"))
ref=TorchDDP(copy.deepcopy(base),device_ids=[rank]); ref2=TorchDDP(copy.deepcopy(base),device_ids=[rank])
oo=SGD(ours.parameters(),lr=0.002,momentum=0.9,weight_decay=5e-4)
o1=torch.optim.SGD(ref.parameters(),lr=0.002,momentum=0.9,weight_decay=5e-4); o2=torch.optim.SGD(ref2.parameters(),lr=0.002,momentum=0.9,weight_decay=5e-4)
g=torch.Generator(device=dev).manual_seed(7+rank)
def cos(a,b): return F.cosine_similarity(a.flatten().double(),b.flatten().double(),dim=0).item()
names=[n for n,_ in ours.module.named_parameters()]
for it in range(4):
    x=torch.randn(16,3,32,32,device=dev,generator=g); t=torch.randint(0,10,(16,),device=dev,generator=g)
    for m,o in ((ours,oo),(ref,o1),(ref2,o2)):
        o.zero_grad(set_to_none=True); F.cross_entropy(m(x),t).backward()
    torch.cuda.synchronize()
    if rank==0:
        worst=(1,None); worst2=(1,None)
        for n,a,b,c in zip(names,ours.module.parameters(),ref.module.parameters(),ref2.module.parameters()):
            c1=cos(a.grad,b.grad); c2=cos(b.grad,c.grad)
            if c1<worst[0]: worst=(c1,n)
            if c2<worst2[0]: worst2=(c2,n)
        print(f"step {it}: worst grad cos ours-vs-torch {worst}  torch-vs-torch {worst2}", flush=True)
    if rank==0:
        def rel(a,b): return ((a-b).abs().max()/(b.abs().max()+1e-30)).item()
        print("   grad relerr worst:", max((rel(a.grad,b.grad),n) for n,a,b in zip(names,ours.module.parameters(),ref.module.parameters())), flush=True)
    for o in (oo,o1,o2): o.step()
    torch.cuda.synchronize()
    if rank==0:
        print("   param relerr worst:", max((rel(a,b),n) for n,a,b in zip(names,ours.module.parameters(),ref.module.parameters())))
        st=o1.state_dict()["state"]; so=oo.state_dict()["state"]
        print("   momentum relerr worst:", max((rel(so[i]["momentum_buffer"],st[i]["momentum_buffer"]),names[i]) for i in range(len(names))), flush=True)
torch.cuda.synchronize()
if rank==0:
    for n,a,b,c in zip(names,ours.module.parameters(),ref.module.parameters(),ref2.module.parameters()):
        c1=cos(a,b); c2=cos(b,c)
        if c1<0.99999 or c2<0.99999: print(f"param {n}: ours-vs-torch {c1:.6f} torch-vs-torch {c2:.6f}")
ours.shutdown(); dist.destroy_process_group()
