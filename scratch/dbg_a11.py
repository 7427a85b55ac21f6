import torch, torch.nn.functional as F, sys
sys.path.insert(0,'.')
from oracle import quant
from vllm_xpu_kernels_b200 import ops
E4=torch.float8_e4m3fn
for dtype in (torch.bfloat16, torch.float16):
  for qd in (E4, torch.int8):
    for (T,H,G) in ((1,1024,64),(16,5120,128)):
        torch.manual_seed(0)
        x=torch.randn(T,2*H,dtype=dtype)*(1.0/H)
        rq,rs=quant.silu_and_mul_per_block_quant(x,G,qd)
        q,s=ops.silu_and_mul_per_block_quant(x.cuda(),G,qd)
        s=s.cpu(); q8=q.cpu().view(torch.uint8); r8=rq.view(torch.uint8)
        y=(F.silu(x[:,:H])*x[:,H:]).float()
        neq=(q8!=r8)
        print(dtype,qd,T,H,G,"scale mismatch",(s!=rs).float().mean().item(),"code mismatch",neq.float().mean().item())
        idx=neq.nonzero()[:6]
        for (t,h) in idx.tolist():
            g=h//G
            print("   t,h",t,h,"y",y[t,h].item(),"s_ref",rs[t,g].item(),"s_gpu",s[t,g].item(),"y/s",(y[t,h]/rs[t,g]).item(),"ref code",r8[t,h].item(),"gpu code",q8[t,h].item())
