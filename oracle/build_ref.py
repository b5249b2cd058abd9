"""TEST / BASELINE INFRASTRUCTURE — builds oracle/_ref/libref_attn.so from the REFERENCE's own sources.

Compiles /root/reference/lightglue_attention_plugin/attention_headdim_64_fp16in_{fp16,fp32}out.cu
*where they lie* (nothing is copied into this repo) for sm_100a, with the reference's vendored
CUTLASS 3.4.1 headers and an empty `torch/torch.h` stub (the reference includes it but uses no
torch symbol — SURVEY.md §0.1), plus oracle/ref_shim.cu.  Flags follow the reference's
CMakeLists.txt:11-12 (`-O3 -DNDEBUG -w --use_fast_math`) except the arch.  Outputs only under
oracle/_ref/ (git-ignored, travels to the GPU box).  No-op when /root/reference is absent.
"""
from __future__ import annotations

import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
LIB = OUT / "libref_attn.so"
REF = Path("/root/reference")
PLUG = REF / "lightglue_attention_plugin"
CUTLASS = REF / "third_party" / "cutlass-v3.4.1" / "include"


def build(force: bool = False) -> Path | None:
    if not PLUG.exists() or not CUTLASS.exists():
        return LIB if LIB.exists() else None
    if LIB.exists() and not force and LIB.stat().st_mtime >= (HERE / "ref_shim.cu").stat().st_mtime:
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    stub = OUT / "stub" / "torch"
    stub.mkdir(parents=True, exist_ok=True)
    (stub / "torch.h").write_text("// empty stub: the reference includes <torch/torch.h> but uses nothing from it\n")
    common = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-DNDEBUG", "-w", "--use_fast_math",
              "-std=c++17", "-Xcompiler", "-fPIC", "-I", str(CUTLASS), "-I", str(OUT / "stub"), "-I", str(PLUG)]
    srcs = [PLUG / "attention_headdim_64_fp16in_fp16out.cu", PLUG / "attention_headdim_64_fp16in_fp32out.cu",
            HERE / "ref_shim.cu"]
    objs = [OUT / (s.stem + ".o") for s in srcs]

    def cc(pair):
        s, o = pair
        r = subprocess.run(common + ["-c", str(s), "-o", str(o)], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s}:\n{r.stdout}{r.stderr}")

    with ThreadPoolExecutor(3) as ex:
        list(ex.map(cc, zip(srcs, objs)))
    r = subprocess.run([nvcc, "-shared", "-cudart", "shared", "-o", str(LIB)] + [str(o) for o in objs],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    for o in objs:
        o.unlink()
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
