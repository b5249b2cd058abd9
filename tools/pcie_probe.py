"""GPU box: pinned-host <-> device copy rates by chunk size, H2D alone and with D2H running beside it (the e2e leg's ceiling).

    python tools/pcie_probe.py
"""
import json
import time

import torch

dev = torch.device("cuda", 0)
TOTAL = 192 << 20
src = torch.empty(TOTAL, dtype=torch.uint8).pin_memory()
dst = torch.empty(TOTAL // 3, dtype=torch.uint8).pin_memory()
d_in = torch.empty(TOTAL, dtype=torch.uint8, device=dev)
d_out = torch.empty(TOTAL // 3, dtype=torch.uint8, device=dev)
s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def run(chunk, with_d2h):
    n = TOTAL // chunk
    oc = chunk // 3
    def once():
        for i in range(n):
            with torch.cuda.stream(s_in):
                d_in[i * chunk:(i + 1) * chunk].copy_(src[i * chunk:(i + 1) * chunk], non_blocking=True)
            if with_d2h:
                with torch.cuda.stream(s_out):
                    dst[i * oc:(i + 1) * oc].copy_(d_out[i * oc:(i + 1) * oc], non_blocking=True)
        torch.cuda.synchronize(dev)
    once()
    t0 = time.perf_counter()
    for _ in range(5):
        once()
    dt = (time.perf_counter() - t0) / 5
    return round(TOTAL / dt / 1e9, 1), round((TOTAL // 3 if with_d2h else 0) / dt / 1e9, 1)


for chunk in (3 << 18, 3 << 19, 3 << 20, 3 << 21, 3 << 23, 3 << 25):
    a = run(chunk, False)
    b = run(chunk, True)
    print(json.dumps({"h2d_chunk_KiB": chunk >> 10, "h2d_alone_GBps": a[0], "h2d_with_d2h_GBps": b[0], "d2h_beside_GBps": b[1]}), flush=True)
