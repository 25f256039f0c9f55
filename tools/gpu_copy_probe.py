import torch, time
dev = torch.device("cuda:0")
img = torch.randint(0, 256, (3, 4096, 4096), dtype=torch.uint8)
d = img.to(dev); torch.cuda.synchronize()
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("H2D pageable .to():        %.1f ms" % t(lambda: img.to(dev)))
print("H2D pin_memory() + to():   %.1f ms" % t(lambda: img.pin_memory().to(dev, non_blocking=True)))
print("D2H .cpu():                %.1f ms" % t(lambda: d.cpu()))
def pinned_out():
    r = torch.empty(d.shape, dtype=torch.uint8, pin_memory=True); r.copy_(d, non_blocking=True); torch.cuda.current_stream().synchronize(); return r
print("D2H into fresh pinned:     %.1f ms" % t(pinned_out))
keep = [pinned_out() for _ in range(3)]
print("D2H into fresh pinned (no reuse of freed blocks): %.1f ms" % t(lambda: keep.append(pinned_out())))
