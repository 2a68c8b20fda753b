import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_gpu_lstm import _setup, _px
from dinunet_implementations_b200.ops.lstm import lstm_forward, lstm_backward
def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/n*1000
import os
for nbe in ('4','8','16'):
  os.environ['DNB_LSTM_NBE']=nbe
  for (B,S,H,D) in [(16,98,192,2),(16,196,192,2),(32,98,192,2)]:
      lstm, enc = _setup(B,S,H,D,E=64)
      with torch.no_grad():
          whh=[c.h2h.weight.detach().to(torch.bfloat16).contiguous() for c in lstm.lstms]; bhh=[c.h2h.bias.detach().contiguous() for c in lstm.lstms]
          px=_px(lstm, enc, B,S,H,D)
          Hmean,hT,cT,C,Hprev=lstm_forward(px.clone(),whh,bhh,B,S,H,D)
          db=[torch.zeros(4*H,device='cuda') for _ in range(D)]
          dH=torch.randn(B,D*H,device='cuda')
          pxs=[px.clone() for _ in range(60)]
          it=iter(range(10**9))
          tf=timeit(lambda: lstm_forward(pxs[next(it)%60],whh,bhh,B,S,H,D))
          tb=timeit(lambda: lstm_backward(px,C,whh,dH,B,S,H,D,db))
          print("nbe",nbe,(B,S,H,D),"fwd us %.1f  bwd us %.1f"%(tf,tb))
