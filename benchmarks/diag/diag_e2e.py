import sys, time, torch
sys.path.insert(0, '.')
from dinunet_implementations_b200.parallel.engine import DeviceEngine, make_ica_trainer
tr = make_ica_trainer(0, batch_size=16, precision_bits="16")
eng = DeviceEngine(tr, use_graph=True)
g = torch.Generator().manual_seed(0)
host = [{"inputs": torch.randn(16, 98, 100, 10, generator=g).pin_memory(), "labels": torch.randint(0, 2, (16,), generator=g).pin_memory()} for _ in range(8)]
for i in range(10): eng.train_step_async(host[i % 8])
eng.drain(); torch.cuda.synchronize()
N = 400
t0 = time.perf_counter()
for i in range(N): eng.train_step_async(host[i % 8])
t1 = time.perf_counter()
eng.drain(); torch.cuda.synchronize()
t2 = time.perf_counter()
print("host issue us/step %.1f   total us/step %.1f" % ((t1 - t0) / N * 1e6, (t2 - t0) / N * 1e6))
dev = [eng.to_device(h) for h in host]
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(N): eng.train_step(dev[i % 8])
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("device-batch: host issue us/step %.1f   total us/step %.1f" % ((t1 - t0) / N * 1e6, (t2 - t0) / N * 1e6))
