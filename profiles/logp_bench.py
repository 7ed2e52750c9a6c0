import torch, sys
sys.path.insert(0, ".")
from glowtts_b200 import ops
B,D,TX,TY=32,80,180,860
g=torch.Generator(device="cuda").manual_seed(1)
xm=torch.randn(B,D,TX,device="cuda",generator=g); z=torch.randn(B,D,TY,device="cuda",generator=g)
xl=0.3*torch.randn(B,D,TX,device="cuda",generator=g)
for logs in (None, xl):
    for _ in range(3): out=ops.logp(xm,logs,z)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): out=ops.logp(xm,logs,z)
    e1.record(); torch.cuda.synchronize()
    l = torch.zeros_like(xm) if logs is None else logs
    s = torch.exp(-2*l.double())
    ref = (-0.5*torch.log(torch.tensor(2*torch.pi, dtype=torch.float64)) - l.double() - 0.5*xm.double()**2*s).sum(1).unsqueeze(-1) \
        + torch.matmul((xm.double()*s).transpose(1,2), z.double()) + torch.matmul((-0.5*s).transpose(1,2), z.double()**2)
    print("logs" if logs is not None else "mean_only", "us", round(e0.elapsed_time(e1)/20*1e3,1), "max abs err", float((out.double()-ref).abs().max()), "max |logp|", float(ref.abs().max()))
