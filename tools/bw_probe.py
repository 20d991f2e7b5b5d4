import torch
def t(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3
for mb in (26, 52, 105, 210, 1024):
    n = mb * (1 << 20) // 4
    a = torch.empty(n, device="cuda"); b = torch.empty(n, device="cuda")
    tf = t(lambda: a.fill_(1.0)); tc = t(lambda: b.copy_(a)); tr = t(lambda: a.sum())
    print("%5d MB: fill %.2f TB/s  copy(r+w) %.2f TB/s  read(sum) %.2f TB/s" % (mb, n*4/tf/1e12, 2*n*4/tc/1e12, n*4/tr/1e12))
