import sys, torch
sys.path.insert(0, '.')
from dinunet_implementations_b200.ops.gemm import gemm_bf16
torch.manual_seed(0)
def timeit(fn, n=200):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): fn()
    g.replay(); torch.cuda.synchronize()
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n//20): g.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/n*1000
# (M,N,K, kind)
cases = [("enc   relu bf16 +T", 1568,256,1000, dict(out_dtype=torch.bfloat16, relu=True, tr=True)),
         ("i2h   fp32 bias   ", 1568,1536,256, dict(bias=True)),
         ("d_enc mask colsumT", 1568,256,1536, dict(mask=True, colsum=True, tr=True, want_out=False)),
         ("dW_ih split acc   ", 1536,256,1568, dict(acc=True)),
         ("dW_hh split acc   ", 768,192,1568, dict(acc=True)),
         ("dWenc split acc   ", 256,1000,1568, dict(acc=True)),
         ("tiny              ", 128,32,64, dict())]
for name,M,N,K,o in cases:
    a=(torch.randn(M,K,device='cuda')*0.5).to(torch.bfloat16); b=(torch.randn(N,K,device='cuda')*0.5).to(torch.bfloat16)
    for bn in (0,32,64,128,256):
        if bn > 32 and N < bn//2: continue
        kw=dict(block_n=bn)
        out=None
        if o.get('acc'):
            out=torch.zeros(M,N,device='cuda'); kw.update(out=out, accumulate=True, split_k=0)
        if o.get('out_dtype'): kw['out_dtype']=o['out_dtype']
        if o.get('relu'): kw['relu']=True
        if o.get('bias'): kw['bias_n']=torch.randn(N,device='cuda')
        if o.get('tr'):
            kw['out_t']=torch.empty(N,(M+7)//8*8,device='cuda',dtype=torch.bfloat16)[:, :M]
        if o.get('mask'): kw['mask']=(torch.randn(M,N,device='cuda')).to(torch.bfloat16)
        if o.get('colsum'): kw['colsum']=torch.zeros(N,device='cuda')
        if o.get('want_out') is False: kw['want_out']=False
        try:
            t=timeit(lambda: gemm_bf16(a,b,**kw))
            print(f"{name} M{M} N{N} K{K} bn{bn:3d}: {t:7.2f} us", flush=True)
        except Exception as e:
            print(name,bn,'EXC',str(e)[:100], flush=True)
