import os, sys, torch
sys.path.insert(0, "/root/repo")
from matcha_b200._lib import Plan
from matcha_b200 import synthetic as weights
def rel(a,b): return float((a.double()-b.double()).norm()/b.double().norm())
sd = weights.estimator_state_dict(160, seed=5)
for prec in ("fp16","bf16"):
  for (b,t,lengths) in [(2,136,[136,77]),(5,452,[451,452,300,9,200])]:
    g = torch.Generator().manual_seed(3)
    x, mu = torch.randn(b,80,t,generator=g), torch.randn(b,80,t,generator=g)
    mask = weights.prefix_mask(lengths,t)
    outs=[]
    for fuse in ("1","0","0","1"):
        os.environ["MB200_FUSE_FF"]=fuse
        plan = Plan(prec,160,use_graph=False); plan.load_estimator(sd)
        outs.append(plan.estimator_forward(x,mask,mu,torch.tensor(0.4)).cpu())
    os.environ["MB200_FUSE_FF"]="0"
    p32 = Plan("fp32",160,use_graph=False); p32.load_estimator(sd)
    ref = p32.estimator_forward(x,mask,mu,torch.tensor(0.4)).cpu()
    print(prec,b,t,"fused-vs-unfused",rel(outs[0],outs[1]),"unfused-vs-unfused",rel(outs[1],outs[2]),"fused-vs-fused",rel(outs[0],outs[3]),
          "fused-vs-fp32",rel(outs[0],ref),"unfused-vs-fp32",rel(outs[1],ref), flush=True)
