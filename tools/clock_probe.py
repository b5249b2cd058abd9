"""GPU box: run the batched attention shape back to back for a few seconds and sample SM clock /
power / throttle reasons (is the kernel power-limited?).  python tools/clock_probe.py [seconds]"""
import subprocess, sys, threading, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
Q = "clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.hw_thermal_slowdown"
proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", "0"], stdout=subprocess.PIPE, text=True)
samples = []
threading.Thread(target=lambda: [samples.append((time.time(), l.strip())) for l in proc.stdout], daemon=True).start()
b, n = 64, 2048
q = torch.randn(b, 4, n, 64, device="cuda").half(); k = torch.randn_like(q); v = torch.randn_like(q); o = torch.empty_like(q)
time.sleep(0.5)
for tag, fn in (("ours", lambda: pkg.attention(q, k, v, out=o)), ("sdpa", lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v))):
    torch.cuda.synchronize(); t0 = time.time(); iters = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(50): fn()
        iters += 50
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize(); t1 = time.time()
    us = e0.elapsed_time(e1) * 1e3 / iters
    mine = [s for (t, s) in samples if t0 + 0.3 < t < t1]
    print(f"{tag}: {us:.1f} us/launch  {4.0*b*4*n*n*64/us/1e6:.0f} TFLOP/s over {secs}s; samples (sm MHz, mem MHz, W, C, pwr_cap, hw_slow, sw_therm, hw_therm):")
    for s in mine[:: max(1, len(mine) // 8)]: print("   ", s)
    time.sleep(1.0)
proc.terminate()
