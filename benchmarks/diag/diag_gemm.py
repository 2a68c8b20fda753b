import sys, torch
sys.path.insert(0, '.')
from dinunet_implementations_b200.ops.gemm import gemm_bf16
torch.manual_seed(0)
for (M,N,K,bn) in [(128,16,64,16),(128,32,64,32),(128,64,64,64),(128,128,64,128),(128,128,256,128),(256,256,512,0),(1568,256,1000,0)]:
    a=(torch.randn(M,K,device='cuda')*0.5).to(torch.bfloat16); b=(torch.randn(N,K,device='cuda')*0.5).to(torch.bfloat16)
    try:
        out=gemm_bf16(a,b,block_n=bn); torch.cuda.synchronize()
        ref=a.float()@b.float().t()
        print(M,N,K,bn,'maxerr',(out-ref).abs().max().item(),'refmax',ref.abs().max().item(), 'row0', out[0,:4].tolist(), ref[0,:4].tolist(), flush=True)
    except Exception as e:
        print(M,N,K,bn,'EXC',e, flush=True); break
