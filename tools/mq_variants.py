"""Build / time named -D variants of the library (tuning aid for the multi-Q-tile kernel).

    python tools/mq_variants.py build mw1:-DLG_ATTN_MQ_MMA_WARPS=1 mw4:-DLG_ATTN_MQ_MMA_WARPS=4,-DFOO=1   # here
    python tools/mq_variants.py run --modes 2,3 mw1 mw4                                                    # GPU box
"""
import importlib.util
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"
VDIR = CSRC / "build" / "variants"


def lib_of(tag):
    return VDIR / tag / "liblightglue_attention_plugin.so"


if sys.argv[1] == "build":
    spec = importlib.util.spec_from_file_location("lgattn_build", CSRC / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    def one(arg):
        tag, _, defs = arg.partition(":")
        mod.build(force=True, defs=[d for d in defs.split(",") if d], out=lib_of(tag))
        return tag

    with ThreadPoolExecutor(8) as ex:
        for t in ex.map(one, sys.argv[2:]):
            print("built", t, flush=True)
else:
    args = sys.argv[2:]
    modes = "2,3"
    if args and args[0] == "--modes":
        modes, args = args[1], args[2:]
    for tag in args:
        subprocess.run([sys.executable, str(ROOT / "tools" / "mq_probe.py"), "--modes", modes, "--lib", str(lib_of(tag)), "--tag", tag,
                        "--out", "gpurun_out/mq_variants.json"])
