"""Generate the pair-matcher golden fixture FROM THE UNMODIFIED REFERENCE MODEL (build container only;
/root/reference does not exist on the GPU box).

    python tests/golden/make_golden_matcher.py

Runs `LightGlue(features=None)` of /root/reference/lightglue_pytorch_no_plugin/lightglue.py (seeded
random weights — no pretrained weights are available offline) in FP32 on CPU on
synthetic SuperPoint-shaped features, then the reference's own `filter_matches`, and writes
  matcher_forward.npz — per case: inputs, `lightglue_descriptors_0/1`, `lightglue_scores`
                        (the three engine outputs, export_onnx_with_plugin.py:197), the reference's
                        match list / scores, and a SHA-256 of the state dict the outputs belong to.
The weights themselves are NOT stored: `reference_init_state_dict(seed)` of the product regenerates
them bit-for-bit (same torch CPU RNG stream and initialisers), which the SHA-256 verifies.
"""
from __future__ import annotations

import hashlib
import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
REF = Path("/root/reference")
sys.path.insert(0, str(ROOT))

# (name, weight seed, n_layers, feature seed, n0, n1)
CASES = [
    ("l9_n160_m130", 0, 9, 1000, 160, 130),
    ("l2_n64_m64", 3, 2, 1001, 64, 64),
    ("l3_n1_m70", 5, 3, 1002, 1, 70),
]


def state_dict_sha256(sd: dict) -> str:
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().numpy()).tobytes())
    return h.hexdigest()


def main() -> None:
    sys.path.insert(0, str(REF))
    from lightglue_pytorch_no_plugin import lightglue as ref_lg
    from oracle import matcher_oracle as mo

    torch.set_grad_enabled(False)
    out = {}
    for name, wseed, n_layers, fseed, n0, n1 in CASES:
        torch.manual_seed(wseed)
        model = ref_lg.LightGlue(features=None, n_layers=n_layers).eval()
        k0, k1, d0, d1 = mo.synth_pair(fseed, n0, n1)
        desc0, desc1, scores = model(k0, k1, d0, d1)
        matches, mscores = ref_lg.filter_matches(scores, model.conf.filter_threshold)
        out[f"{name}.meta"] = np.array([wseed, n_layers, fseed, n0, n1], dtype=np.int64)
        out[f"{name}.sha256"] = np.frombuffer(state_dict_sha256(model.state_dict()).encode(), dtype=np.uint8)
        out[f"{name}.desc0"] = desc0.numpy()
        out[f"{name}.desc1"] = desc1.numpy()
        out[f"{name}.scores"] = scores.numpy()
        out[f"{name}.ref_matches"] = matches.numpy().astype(np.int32)
        out[f"{name}.ref_mscores"] = mscores.numpy()
        # random weights give no match above the default 0.1 threshold: also keep the mutual-NN set itself
        matches_t0, mscores_t0 = ref_lg.filter_matches(scores, 0.0)
        out[f"{name}.ref_matches_t0"] = matches_t0.numpy().astype(np.int32)
        out[f"{name}.ref_mscores_t0"] = mscores_t0.numpy()
        print(name, "matches:", matches.shape[0], "mutual:", matches_t0.shape[0], "scores range", float(scores.min()), float(scores.max()))
    np.savez_compressed(HERE / "matcher_forward.npz", **out)
    print("wrote", HERE / "matcher_forward.npz", (HERE / "matcher_forward.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    main()
