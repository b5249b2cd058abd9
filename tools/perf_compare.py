"""Perf table (GPU box): new sm_100a kernel vs the reference's kernels recompiled for sm_100a
(oracle/_ref, through the enqueue-order shim) vs torch SDPA, on the BASELINE shapes.

    python tools/perf_compare.py [--out gpurun_out/perf.json] [--quick]

Latency shapes are timed as a CUDA-graph chain of CALLS independent problems (the deployment mode,
demo/lightglue_trt.cpp:347-366) whose footprint exceeds L2; batched shapes as back-to-back launches.
"""
import argparse
import ctypes
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from lgattn_pkg import pkg  # noqa: E402

H, D = 4, 64


def load_ref():
    path = ROOT / "oracle" / "_ref" / "libref_attn.so"
    if not path.exists():
        return None
    lib = ctypes.CDLL(str(path))
    lib.ref_attn_workspace_bytes.restype = ctypes.c_size_t
    lib.ref_attn_workspace_bytes.argtypes = [ctypes.c_int] * 4
    lib.ref_attn_enqueue.restype = ctypes.c_int
    lib.ref_attn_enqueue.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_void_p] * 2
    return lib


def time_graph_chain(fn, calls, iters=20, warm=5):
    """fn(i, stream_ptr) enqueues problem i. Returns microseconds per problem."""
    stream = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        for i in range(calls):
            fn(i, stream.cuda_stream)
        stream.synchronize()
        with torch.cuda.graph(graph, stream=stream):
            for i in range(calls):
                fn(i, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for _ in range(warm):
        graph.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * calls)


def time_loop(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    lib = pkg.load()
    ref = load_ref()
    dev = torch.device("cuda:0")
    rows = []

    # ---------------- latency shapes: graph chain of independent problems
    lat_shapes = [(1024, 1024, "f16"), (1024, 1000, "f16"), (777, 1024, "f16"), (1024, 1024, "f32"), (512, 512, "f16"),
                  (2048, 2048, "f16")]
    if args.quick:
        lat_shapes = lat_shapes[:1]
    for (n, m, dt) in lat_shapes:
        calls = 128 if n <= 1024 else 48
        tdt = torch.float16 if dt == "f16" else torch.float32
        q = torch.randn(calls, 1, H, n, D, device=dev).to(tdt)
        k = torch.randn(calls, 1, H, m, D, device=dev).to(tdt)
        v = torch.randn(calls, 1, H, m, D, device=dev).to(tdt)
        o = torch.empty(calls, 1, H, n, D, device=dev, dtype=tdt)
        mode = pkg.F16_F16 if dt == "f16" else pkg.F32_F32
        wsb = lib.lg_attn_hd64_workspace_bytes(1, H, n, m, mode)
        ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=dev)

        def ours(i, s):
            rc = lib.lg_attn_hd64_launch(q[i].data_ptr(), k[i].data_ptr(), v[i].data_ptr(), o[i].data_ptr(), 1, H, n, m,
                                         mode, ws.data_ptr() if wsb else None, wsb, s)
            assert rc == 0, rc

        row = {"shape": f"B1_N{n}_M{m}_{dt}", "flops": 4.0 * H * n * m * D, "ours_us": time_graph_chain(ours, calls)}
        if ref is not None:
            rws = torch.zeros(ref.ref_attn_workspace_bytes(1, H, n, m), dtype=torch.uint8, device=dev)

            def theirs(i, s):
                r = ref.ref_attn_enqueue(q[i].data_ptr(), k[i].data_ptr(), v[i].data_ptr(), o[i].data_ptr(), 1, H, n, m,
                                         0 if dt == "f16" else 1, rws.data_ptr(), s)
                assert r > 0

            row["ref_us"] = time_graph_chain(theirs, calls)
        if dt == "f16":
            def sdpa(i, s):
                torch.nn.functional.scaled_dot_product_attention(q[i], k[i], v[i])
            try:
                row["sdpa_us"] = time_graph_chain(sdpa, calls)
            except Exception as e:  # noqa: BLE001
                row["sdpa_us"] = None
                row["sdpa_err"] = str(e)[:80]
        row["ours_tflops"] = row["flops"] / row["ours_us"] / 1e6
        rows.append(row)
        print(json.dumps(row), flush=True)
        del q, k, v, o

    # ---------------- batched shapes
    bat_shapes = [(64, 2048, 2048, "f16"), (64, 2048, 2048, "f32out"), (256, 2048, 2048, "f16"), (16, 4096, 4096, "f16"),
                  (64, 4096, 4096, "f16"), (64, 2000, 1900, "f16"), (128, 1024, 1024, "f16")]
    if args.quick:
        bat_shapes = bat_shapes[:1]
    for (b, n, m, dt) in bat_shapes:
        q = torch.randn(b, H, n, D, device=dev).half()
        k = torch.randn(b, H, m, D, device=dev).half()
        v = torch.randn(b, H, m, D, device=dev).half()
        odt = torch.float32 if dt == "f32out" else torch.float16
        o = torch.empty(b, H, n, D, device=dev, dtype=odt)
        row = {"shape": f"B{b}_N{n}_M{m}_{dt}", "flops": 4.0 * b * H * n * m * D}
        row["ours_us"] = time_loop(lambda: pkg.attention(q, k, v, out_dtype=odt, out=o))
        row["ours_tflops"] = row["flops"] / row["ours_us"] / 1e6
        if ref is not None and dt == "f16" and b * n <= 64 * 4096:
            rws = torch.zeros(ref.ref_attn_workspace_bytes(b, H, n, m), dtype=torch.uint8, device=dev)
            s = torch.cuda.current_stream().cuda_stream
            row["ref_us"] = time_loop(lambda: ref.ref_attn_enqueue(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), b, H,
                                                                   n, m, 0, rws.data_ptr(), s))
            row["ref_tflops"] = row["flops"] / row["ref_us"] / 1e6
            del rws
        if dt == "f16":
            row["sdpa_us"] = time_loop(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v))
            row["sdpa_tflops"] = row["flops"] / row["sdpa_us"] / 1e6
        rows.append(row)
        print(json.dumps(row), flush=True)
        del q, k, v, o
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
