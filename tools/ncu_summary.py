"""Condense an .ncu-rep (read here, no GPU) into the handful of numbers the kernel notes quote.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--source N] [--json profiles/x.json]   # N hottest SASS lines by stall samples
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__cycles_elapsed.max.per_second", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"]


def raw(rep, json_out=None):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for n, vals in enumerate(rows[2:]):
        d = dict(zip(hdr, vals))
        if json_out and n == 0:
            import json
            u = dict(zip(hdr, units))
            scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}
            keep = {"kernel": d.get("Kernel Name"), "source_report": rep}
            for k in KEYS:
                if k in d and d[k] != "":
                    keep[k] = float(d[k].replace(",", "")) * (scale.get(u.get(k, ""), 1.0) if k.startswith("dram__bytes") or k.startswith("lts__t_bytes") else 1.0)
                    if not (k.startswith("dram__bytes") or k.startswith("lts__t_bytes")):
                        keep[k + ".unit"] = u.get(k, "")
            stalls_j = {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): float(v)
                        for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v}
            keep["stalls_per_issue"] = dict(sorted(stalls_j.items(), key=lambda kv: -kv[1])[:8])
            open(json_out, "w").write(json.dumps(keep, indent=1) + "\n")
        print("==", d.get("Kernel Name", "?")[:100])
        for k in KEYS:
            if k in d:
                print(f"  {k} = {d[k]}")
        stalls = {k: float(v) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and v}
        if not stalls:
            stalls = {k: float(v) for k, v in d.items() if "issue_stalled" in k and "per_issue_active" in k and v}
        for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:10]:
            print(f"  stall {k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')} = {v:.3f}")


def source(rep, top):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    rows = rows[1:]   # first line names the kernel
    hdr = rows[0]
    si = next((i for i, h in enumerate(hdr) if h.strip() == "# Samples" or "Warp Stall Sampling (All" in h), None)
    ii = next((i for i, h in enumerate(hdr) if h.strip() == "Source"), 1)
    ei = next((i for i, h in enumerate(hdr) if "Instructions Executed" == h.strip()), None)
    print("columns:", [h for h in hdr][:12])
    body = [r for r in rows[1:] if len(r) > max(si or 0, ii)]
    tot = sum(float(r[si] or 0) for r in body)
    ranked = sorted(enumerate(body), key=lambda t: -float(t[1][si] or 0))[:top]
    print(f"total samples {tot:.0f}")
    for idx, r in sorted(ranked):
        print(f"  {idx:5d} {float(r[si] or 0) / tot * 100:5.2f}%  exec={r[ei] if ei else ''}  {r[ii][:110]}")


if __name__ == "__main__":
    rep = sys.argv[1]
    raw(rep, sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None)
    if "--source" in sys.argv:
        source(rep, int(sys.argv[sys.argv.index("--source") + 1]))
