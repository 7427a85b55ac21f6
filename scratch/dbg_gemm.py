import sys, torch, ctypes
sys.path.insert(0,'.')
from oracle import gemm as og, quant as oq, formats as F
from vllm_xpu_kernels_b200 import ops, _loader
lib=_loader.cdll()
E4=torch.float8_e4m3fn
def stats():
    a=ctypes.c_uint64(); b=ctypes.c_uint64(); lib.vxk_gemm_stats(ctypes.byref(a),ctypes.byref(b)); return a.value,b.value
def report(name,out,ref):
    d=(out.float().cpu()-ref.float()).abs()
    rel=d/(ref.float().abs()+1e-3)
    print(f"{name}: max abs {d.max():.4g} max rel {rel.max():.4g} mean abs {d.mean():.4g} ref absmean {ref.float().abs().mean():.4g} stats {stats()}", flush=True)
torch.manual_seed(0)
which=sys.argv[1] if len(sys.argv)>1 else "all"
for (M,N,K) in ((128,128,512),(100,200,2048),(1024,2048,2048),(2048,4224,1024),(1500,3000,1536)):
    a=torch.randn(M,K,dtype=torch.bfloat16)/10; w=torch.randn(N,K,dtype=torch.bfloat16)/10
    if which in ("all","f8"):
        aq,asc=oq.dynamic_per_token_scaled_fp8_quant(a,E4); wq,wsc=oq.dynamic_per_token_scaled_fp8_quant(w,E4)
        ref=og.fp8_gemm_ref(aq,wq.t(),torch.bfloat16,asc,wsc)
        out=ops.fp8_gemm(aq.cuda(),wq.cuda().t(),torch.bfloat16,asc.cuda(),wsc.cuda(),None); torch.cuda.synchronize()
        report(f"f8 per-token/channel {M}x{N}x{K}",out,ref)
    if which in ("all","mx8"):
        sa,qa=F.mx_quantize(a,"mxfp8"); sw,qw=F.mx_quantize(w,"mxfp8")
        ref=og.fp8_gemm_ref(qa,qw.t(),torch.bfloat16,sa,sw.t().contiguous())
        out=ops.fp8_gemm(qa.cuda(),qw.cuda().t(),torch.bfloat16,sa.cuda(),sw.t().contiguous().cuda(),None); torch.cuda.synchronize()
        report(f"mxfp8 {M}x{N}x{K}",out,ref)
    if which in ("all","mx4") and K%512==0:
        a4=a*0.1; w4=w*0.1
        sa,qa=F.mx_quantize(a4,"mxfp4"); sw,qw=F.mx_quantize(w4,"mxfp4")
        ref=og.fp4_gemm_ref(qa,qw.t(),sa,sw,torch.bfloat16)
        out=ops.fp4_gemm(qa.cuda(),qw.cuda().t(),sa.cuda(),sw.cuda(),torch.bfloat16,None); torch.cuda.synchronize()
        report(f"mxfp4 {M}x{N}x{K}",out,ref)
