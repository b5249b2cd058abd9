"""Generate the golden fixtures in this directory FROM THE UNMODIFIED REFERENCE (run in the build
container only; /root/reference does not exist on the GPU box).

    python tests/golden/make_golden.py

Writes
  attn_forward.npz   — `Attention.forward` of /root/reference/lightglue_pytorch_no_plugin/lightglue.py
                       (:75-85) on seeded FP16-rounded Q/K/V at small ragged shapes; outputs FP32.
  attn_sdpa.npz      — `MHAHeadDim64.forward` (eager twin of the plugin,
                       lightglue_pytorch_with_plugin/lightglue.py:35-46) on the same inputs.
  callsites.npz      — Q/K/V and outputs captured at attention call sites of a seeded
                       `LightGlue(features=None)` forward (layers 0/8, self and both cross
                       directions; lightglue.py:148,194,198) on synthetic SuperPoint-shaped features.
The reference publishes no tests / vectors of its own, so these reference-generated
outputs are what pins oracle/attention_oracle.py.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference")

# (seed, B, H, N, M): small, ragged, non-tile-multiple, N != M, N or M == 1
SHAPES = [
    (11, 1, 4, 64, 64),
    (12, 1, 4, 37, 51),
    (13, 1, 4, 130, 200),
    (14, 1, 4, 1, 77),
    (15, 1, 4, 65, 1),
    (16, 2, 4, 96, 129),
    (17, 1, 4, 129, 127),
]


def synth_features(seed: int, n: int):
    """SuperPoint-shaped synthetic features (SURVEY.md §8d "D-feat")."""
    g = torch.Generator().manual_seed(seed)
    xs = torch.randint(4, 636, (n,), generator=g).float()
    ys = torch.randint(4, 476, (n,), generator=g).float()
    kpts = (torch.stack([xs, ys], -1) - torch.tensor([320.0, 240.0])) / 320.0
    desc = torch.randn(n, 256, generator=g)
    desc = desc / desc.norm(dim=-1, keepdim=True)
    return kpts[None], desc[None]


def main() -> None:
    if not REF.exists():
        sys.exit("reference tree not present: fixtures can only be regenerated in the build container")
    sys.path.insert(0, str(REF))
    torch.set_grad_enabled(False)
    from lightglue_pytorch_no_plugin import lightglue as lg_no
    from lightglue_pytorch_with_plugin import lightglue as lg_with

    sys.path.insert(0, str(HERE.parent.parent))
    from oracle.attention_oracle import make_qkv

    fwd, sdpa = {}, {}
    attn = lg_no.Attention()
    for seed, b, h, n, m in SHAPES:
        q, k, v = make_qkv(seed, b, h, n, m)
        tq, tk, tv = (torch.from_numpy(x.astype(np.float32)) for x in (q, k, v))
        tag = f"s{seed}_b{b}_h{h}_n{n}_m{m}"
        fwd[tag + "_q"], fwd[tag + "_k"], fwd[tag + "_v"] = q, k, v
        fwd[tag + "_o"] = attn(tq, tk, tv).numpy()
        sdpa[tag + "_o"] = lg_with.MHAHeadDim64.forward(None, tq, tk, tv).numpy()
    np.savez_compressed(HERE / "attn_forward.npz", **fwd)
    np.savez_compressed(HERE / "attn_sdpa.npz", **sdpa)

    # call-site capture from the reference model with seeded random weights
    torch.manual_seed(0)
    model = lg_no.LightGlue(features=None).eval()
    captured = []
    orig = lg_no.Attention.forward

    def recording_forward(self, query, key, value):
        out = orig(self, query, key, value)
        captured.append((query.clone(), key.clone(), value.clone(), out.clone()))
        return out

    lg_no.Attention.forward = recording_forward
    k0, d0 = synth_features(1000, 64)
    k1, d1 = synth_features(1001, 48)
    outs = model(k0, k1, d0, d1)
    lg_no.Attention.forward = orig
    assert len(captured) == 36, len(captured)   # 9 layers x (self0, self1, cross01, cross10)
    cs = {}
    for layer in (0, 8):
        for j, name in enumerate(("self0", "self1", "cross01", "cross10")):
            q, k, v, o = captured[layer * 4 + j]
            tag = f"L{layer}_{name}"
            cs[tag + "_q"], cs[tag + "_k"], cs[tag + "_v"], cs[tag + "_o"] = (
                q.numpy(), k.numpy(), v.numpy(), o.numpy())
    cs["scores"] = outs[2].numpy() if isinstance(outs, (tuple, list)) else np.zeros(1)
    np.savez_compressed(HERE / "callsites.npz", **cs)
    for f in ("attn_forward.npz", "attn_sdpa.npz", "callsites.npz"):
        print(f, (HERE / f).stat().st_size, "bytes")


if __name__ == "__main__":
    main()
