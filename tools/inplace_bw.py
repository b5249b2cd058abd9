import torch
def timeit(fn, iters=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters
for rows in (32768, 131072, 524288):
    h = torch.randn(rows, 512, device="cuda").half()
    g = torch.empty_like(h)
    us = timeit(lambda: h.mul_(1.0001))
    print(f"in-place mul rows={rows}: {us:7.1f} us {rows*2048/us/1e3:7.1f} GB/s")
    us = timeit(lambda: torch.mul(h, 1.0001, out=g))
    print(f"out-of-place mul rows={rows}: {us:7.1f} us {rows*2048/us/1e3:7.1f} GB/s")
    us = timeit(lambda: g.copy_(h))
    print(f"copy rows={rows}: {us:7.1f} us {rows*2048/us/1e3:7.1f} GB/s")
    us = timeit(lambda: torch.nn.functional.gelu(h))
    print(f"torch gelu (erf) rows={rows}: {us:7.1f} us {rows*2048/us/1e3:7.1f} GB/s")
