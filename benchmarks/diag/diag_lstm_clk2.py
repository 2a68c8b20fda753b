import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_gpu_lstm import _setup, _px
from dinunet_implementations_b200.ops.lstm import lstm_forward, lstm_backward
for (B,S,H,D) in [(16,98,32,1),(16,98,128,1),(16,98,192,2),(16,98,192,2),(16,98,256,1),(64,98,192,2)]:
    lstm, enc = _setup(B,S,H,D,E=64)
    with torch.no_grad():
        whh=[c.h2h.weight.detach().to(torch.bfloat16).contiguous() for c in lstm.lstms]; bhh=[c.h2h.bias.detach().contiguous() for c in lstm.lstms]
        for rep in range(2):
            px=_px(lstm, enc, B,S,H,D)
            dbg=torch.zeros(16,dtype=torch.int64,device='cuda')
            Hmean,hT,cT,C,Hprev=lstm_forward(px,whh,bhh,B,S,H,D,dbg=dbg)
            db=[torch.zeros(4*H,device='cuda') for _ in range(D)]
            dbg2=torch.zeros(16,dtype=torch.int64,device='cuda')
            dP=lstm_backward(px,C,whh,torch.randn(B,D*H,device='cuda'),B,S,H,D,db,dbg=dbg2,whhT=[w.t().contiguous() for w in whh])
            torch.cuda.synchronize()
        f=dbg.tolist(); b=dbg2.tolist()
        print("   fwd pro/loop/epi", f[13]-f[12], f[14]-f[13], f[15]-f[14], " bwd pro/loop/epi", b[13]-b[12], b[14]-b[13], b[15]-b[14])
        print((B,S,H,D),"ncta",H//32,"fwd step",f[11]-f[0],[f[i+1]-f[i] for i in range(11)],"| bwd step",b[11]-b[0],[b[i+1]-b[i] for i in range(11)])
