import sys, json, torch
sys.path.insert(0, "/root/repo")
from lgattn_pkg import pkg
from tools.pairs_sweep import feats
dev = torch.device("cuda:0")
eng = pkg.NativeMatcher(seed=0, device=0)
py = pkg.PairMatcher(seed=0, device=dev)
for (n, b) in [(1024, 1), (1024, 64), (2048, 16), (2048, 32), (4096, 8)]:
    f = feats(b, n, dev)
    f = (f[0].contiguous(), f[1].contiguous(), f[2].contiguous(), f[3].contiguous())
    res = {}
    for name, fn in (("native", lambda: eng.match(*f)), ("python", lambda: py.match(*f))):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        reps = max(2, 128 // b)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = round(reps * b / (e0.elapsed_time(e1) * 1e-3), 1)
    print(json.dumps({"kpts": n, "pairs_per_forward": b, **res}), flush=True)
    py._plans.clear(); torch.cuda.empty_cache()
