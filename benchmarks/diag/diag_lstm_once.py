import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_gpu_lstm import _setup, _px
from dinunet_implementations_b200.ops.lstm import lstm_forward, lstm_backward
B,S,H,D = 16,98,192,2
lstm, enc = _setup(B,S,H,D,E=64)
with torch.no_grad():
    whh=[c.h2h.weight.detach().to(torch.bfloat16).contiguous() for c in lstm.lstms]; bhh=[c.h2h.bias.detach().contiguous() for c in lstm.lstms]
    whhT=[w.t().contiguous() for w in whh]
    px=_px(lstm, enc, B,S,H,D)
    for _ in range(3):
        Hmean,hT,cT,C,Hprev=lstm_forward(px.clone(),whh,bhh,B,S,H,D)
        db=[torch.zeros(4*H,device='cuda') for _ in range(D)]
        dP=lstm_backward(px,C,whh,torch.randn(B,D*H,device='cuda'),B,S,H,D,db,whhT=whhT)
    torch.cuda.synchronize()
print("ok")
