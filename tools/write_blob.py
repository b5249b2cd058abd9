"""Write the native engine's weight blob (include/lg_matcher.h) for `lg_pair_bench`.

    python tools/write_blob.py OUT.blob [--layers 9] [--seed 0] [--checkpoint superpoint_lightglue.pth]
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from lgattn_pkg import pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("out")
ap.add_argument("--layers", type=int, default=9)
ap.add_argument("--seed", type=int, default=0)
ap.add_argument("--checkpoint", default=None)
a = ap.parse_args()
m = pkg.matcher
sd = m.load_checkpoint(a.checkpoint, a.layers) if a.checkpoint else m.reference_init_state_dict(a.seed, a.layers)
blob = m.weights_blob(m.MatcherWeights(sd, a.layers, device="cpu"))
Path(a.out).write_bytes(blob)
print(f"wrote {a.out}: {len(blob)} bytes, {a.layers} layers")
