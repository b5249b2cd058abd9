"""In-tree build of liblightglue_attention_plugin.so for sm_100a (no CMake needed).

    python build.py            # build if sources are newer than the library
    python build.py --force

One shared library carries the CUDA kernel, the C-ABI (include/lg_attn.h) and the TensorRT plugin
TU.  Real TensorRT headers are used when found (TENSORRT_ROOT, /usr/include, /usr/local/tensorrt);
otherwise the plugin compiles against plugin/mock_nvinfer/ and the mock host is linked in.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

CSRC = Path(__file__).resolve().parent
BUILD = CSRC / "build"
LIB = BUILD / "liblightglue_attention_plugin.so"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
SOURCES = [CSRC / "lg_attn_api.cu", CSRC / "lg_match_kernels.cu", CSRC / "lg_ffn_sm100.cu", CSRC / "lg_matcher.cu", CSRC / "plugin" / "lightglue_attention_plugin.cpp",
           CSRC / "plugin" / "mock_trt_host.cpp"]
HEADERS = [CSRC / "tools" / "lg_pair_bench.cpp", *sorted(CSRC.glob("*.cuh")), CSRC / "attn_sm100.h",
           CSRC / "plugin" / "lightglue_attention_plugin.h", CSRC / "plugin" / "mock_nvinfer" / "NvInfer.h",
           CSRC.parent.parent / "include" / "lg_attn.h", CSRC.parent.parent / "include" / "lg_match.h", CSRC.parent.parent / "include" / "lg_matcher.h"]


def find_tensorrt_include() -> Path | None:
    roots = [os.environ.get("TENSORRT_ROOT"), "/usr", "/usr/local/tensorrt", "/usr/local/TensorRT", "/opt/tensorrt"]
    for r in roots:
        if not r:
            continue
        for sub in ("include", "include/x86_64-linux-gnu", "include/aarch64-linux-gnu"):
            p = Path(r) / sub / "NvInfer.h"
            if p.exists():
                return p.parent
    return None


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(exe).exists():
        raise RuntimeError("nvcc not found")
    return exe


def up_to_date() -> bool:
    if not LIB.exists():
        return False
    t = LIB.stat().st_mtime
    return all(p.stat().st_mtime <= t for p in SOURCES + HEADERS + [Path(__file__)])


def build(force: bool = False, verbose: bool = False, defs: list[str] | None = None, out: Path | None = None) -> Path:
    """Default: the product library.  `defs` / `out` build a tuning variant (extra -D flags) elsewhere."""
    if out is None and defs:
        raise ValueError("the product library is built without extra -D flags; pass out=... for a tuning variant")
    if out is None and not force and up_to_date():
        return LIB
    BUILD.mkdir(exist_ok=True)
    lib = out or LIB
    lib.parent.mkdir(parents=True, exist_ok=True)
    trt_inc = find_tensorrt_include()
    inc = ["-I", str(trt_inc)] if trt_inc else ["-I", str(CSRC / "plugin" / "mock_nvinfer")]
    libs = (["-lnvinfer"] if trt_inc else []) + ["-lcublasLt"]
    cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=default",
           "-shared", "-cudart", "shared", *inc, *(defs or []), "-o", str(lib), *map(str, SOURCES), *libs]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed: " + " ".join(cmd))
    if verbose:
        sys.stderr.write(r.stderr)
    if out is None:
        build_pair_bench()
    return lib


PAIR_BENCH = BUILD / "lg_pair_bench"


def build_pair_bench() -> Path:
    """The native multi-GPU pair-matching harness (csrc/tools/lg_pair_bench.cpp), linked against the library."""
    src = CSRC / "tools" / "lg_pair_bench.cpp"
    cmd = [nvcc(), "-O2", "-std=c++17", "-x", "cu", *ARCH, "-cudart", "shared", "-o", str(PAIR_BENCH), str(src),
           "-L", str(BUILD), "-l:" + LIB.name, "-Xlinker", "-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed: " + " ".join(cmd))
    return PAIR_BENCH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
