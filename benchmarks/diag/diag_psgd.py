import sys, torch
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
from dinunet_implementations_b200.distrib import lowrank
from dinunet_implementations_b200.ops.gemm import gemm_bf16
from dinunet_implementations_b200.parallel.lowrank_fused import _bf16_pad_cols, orthogonalize_batch
from dinunet_implementations_b200.ops.elementwise import transpose_bf16
torch.manual_seed(0)
for (n,m,r) in [(3,64,3),(64,40,4),(3,64,2),(8,64,3)]:
    M=torch.randn(n,m,device='cuda')*0.1
    Q=lowrank.psgd_init_q((n,m),r,5,'cuda')
    mb=_bf16_pad_cols(M); qt=_bf16_pad_cols(Q.t().contiguous())
    P=torch.empty(n*r,device='cuda').view(n,r)
    gemm_bf16(mb,qt,out=P)
    Pref=M@Q
    print((n,m,r),'P err',(P-Pref).abs().max().item(),'scale',Pref.abs().max().item())
    Ph=Pref.clone(); orthogonalize_batch([Ph]); Phr=lowrank.orthogonalize(Pref)
    print('  orth err',(Ph-Phr).abs().max().item())
    mbT=transpose_bf16(mb); pt=_bf16_pad_cols(Phr.t().contiguous())
    Qn=torch.empty(m*r,device='cuda').view(m,r)
    gemm_bf16(mbT,pt,out=Qn); Qnref=M.t()@Phr
    print('  Q err',(Qn-Qnref).abs().max().item(),'scale',Qnref.abs().max().item())
    G=torch.empty(n,m,device='cuda')
    gemm_bf16(_bf16_pad_cols(Phr),_bf16_pad_cols(Qnref),out=G); Gref=Phr@Qnref.t()
    print('  Mhat err',(G-Gref).abs().max().item(),'scale',Gref.abs().max().item())
