import torch, time
dev = torch.device("cuda:0")
def timeit(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for mb in (381, 437, 1024):
    n = mb * 1000 * 1000 // 8
    x = torch.empty(n, dtype=torch.float64, device=dev)
    y = torch.empty(n, dtype=torch.float64, device=dev)
    t = timeit(lambda: x.fill_(1.5)); print("fill %d MB: %.1f us  %.0f GB/s" % (mb, t * 1e3, mb / t))
    t = timeit(lambda: y.copy_(x)); print("copy %d MB (x2 traffic): %.1f us  %.0f GB/s" % (mb, t * 1e3, 2 * mb / t))
    t = timeit(lambda: torch.cuda.memset if False else x.zero_()); print("zero %d MB: %.1f us  %.0f GB/s" % (mb, t * 1e3, mb / t))
    t = timeit(lambda: x.sum()); print("sum(read) %d MB: %.1f us %.0f GB/s" % (mb, t * 1e3, mb / t))
