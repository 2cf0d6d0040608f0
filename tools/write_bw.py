import torch
x = torch.empty(2_000_000_000, dtype=torch.float64, device="cuda")   # 16 GB
for f, name in ((lambda: x.zero_(), "zero_ (memset)"), (lambda: x.fill_(1.5), "fill_ (kernel)")):
    f(); torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print("%s: 16 GB in %.3f ms -> %.0f GB/s" % (name, min(ts), 16e9 / min(ts) / 1e6))
y = torch.empty_like(x[:1_000_000_000]); 
ts=[]
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y.copy_(x[:1_000_000_000]); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print("copy 8 GB -> 8 GB: %.3f ms -> %.0f GB/s (read+write)" % (min(ts), 16e9/min(ts)/1e6))
