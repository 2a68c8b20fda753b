import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_gpu_lstm import _setup, _px
from dinunet_implementations_b200.ops.lstm import lstm_forward, lstm_backward
B,S,H,D=16,98,192,2
lstm, enc = _setup(B,S,H,D,E=256)
with torch.no_grad():
    whh=[c.h2h.weight.detach().to(torch.bfloat16).contiguous() for c in lstm.lstms]; bhh=[c.h2h.bias.detach().contiguous() for c in lstm.lstms]
    for rep in range(3):
        px=_px(lstm, enc, B,S,H,D)
        dbg=torch.zeros(16,dtype=torch.int64,device='cuda')
        e0,e1,e2=[torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e0.record()
        Hmean,hT,cT,C,Hprev=lstm_forward(px,whh,bhh,B,S,H,D,dbg=dbg)
        e1.record()
        db=[torch.zeros(4*H,device='cuda') for _ in range(D)]
        dbg2=torch.zeros(16,dtype=torch.int64,device='cuda')
        dP=lstm_backward(px,C,whh,torch.randn(B,D*H,device='cuda'),B,S,H,D,db,dbg=dbg2)
        e2.record(); torch.cuda.synchronize()
        f=dbg.tolist(); b=dbg2.tolist()
        print("fwd us", e0.elapsed_time(e1)*1e3, "per-step cycles", f[9]-f[0], "deltas", [f[i+1]-f[i] for i in range(9)])
        print("bwd us", e1.elapsed_time(e2)*1e3, "per-step cycles", b[9]-b[0], "deltas", [b[i+1]-b[i] for i in range(9)])
