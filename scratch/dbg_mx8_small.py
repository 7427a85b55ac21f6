import sys, torch
sys.path.insert(0,'.')
from oracle import gemm as og, formats as F
from vllm_xpu_kernels_b200 import ops
torch.manual_seed(0)
M,N,K=128,128,128
a=torch.randn(M,K,dtype=torch.bfloat16)/10; w=torch.randn(N,K,dtype=torch.bfloat16)/10
sa,qa=F.mx_quantize(a,"mxfp8"); sw,qw=F.mx_quantize(w,"mxfp8")
out=ops.fp8_gemm(qa.cuda(),qw.cuda().t(),torch.bfloat16,sa.cuda(),sw.t().contiguous().cuda(),None); torch.cuda.synchronize()
ref=og.fp8_gemm_ref(qa,qw.t(),torch.bfloat16,sa,sw.t().contiguous())
print((out.float().cpu()-ref.float()).abs().max())
