"""CPU tests: the matcher oracle (oracle/matcher_oracle.py) against fixtures generated from the
unmodified reference LightGlue, and the product's host-side weight logic."""
import hashlib
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import matcher_oracle as mo

REF = Path("/root/reference")


def _sha(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(np.ascontiguousarray(sd[k].detach().cpu().numpy()).tobytes())
    return h.hexdigest()


def _cases(golden_dir):
    z = np.load(golden_dir / "matcher_forward.npz")
    names = sorted({k.split(".")[0] for k in z.files})
    return z, names


def _pairs(matches0):
    idx = torch.nonzero(matches0[0] >= 0)[:, 0]
    return torch.stack([idx, matches0[0][idx]], -1).numpy().astype(np.int32)


def test_seeded_init_reproduces_reference_weights(pkg, golden_dir):
    """reference_init_state_dict(seed) == LightGlue(features=None).state_dict() under the same seed."""
    z, names = _cases(golden_dir)
    for name in names:
        wseed, n_layers = (int(v) for v in z[f"{name}.meta"][:2])
        sd = pkg.matcher.reference_init_state_dict(wseed, n_layers)
        assert _sha(sd) == bytes(z[f"{name}.sha256"]).decode(), name


def test_oracle_matches_reference_outputs(pkg, golden_dir):
    z, names = _cases(golden_dir)
    for name in names:
        wseed, n_layers, fseed, n0, n1 = (int(v) for v in z[f"{name}.meta"])
        sd = pkg.matcher.reference_init_state_dict(wseed, n_layers)
        k0, k1, d0, d1 = mo.synth_pair(fseed, n0, n1)
        out = mo.forward(sd, k0, k1, d0, d1, n_layers=n_layers, threshold=0.0)
        np.testing.assert_allclose(out["desc0"].numpy(), z[f"{name}.desc0"], rtol=0, atol=2e-5)
        np.testing.assert_allclose(out["desc1"].numpy(), z[f"{name}.desc1"], rtol=0, atol=2e-5)
        np.testing.assert_allclose(out["scores"].numpy(), z[f"{name}.scores"], rtol=0, atol=5e-5)
        # the oracle's matches0 form <-> the reference's [K,2] match list
        assert np.array_equal(_pairs(out["matches0"]), z[f"{name}.ref_matches_t0"]), name
        idx = z[f"{name}.ref_matches_t0"][:, 0]
        np.testing.assert_allclose(out["mscores0"][0].numpy()[idx], z[f"{name}.ref_mscores_t0"], rtol=1e-4, atol=1e-12)
        out01 = mo.forward(sd, k0, k1, d0, d1, n_layers=n_layers, threshold=0.1)
        assert np.array_equal(_pairs(out01["matches0"]), z[f"{name}.ref_matches"].reshape(-1, 2))


def test_oracle_fp64_agrees_with_fp32(pkg):
    sd = pkg.matcher.reference_init_state_dict(7, 2)
    k0, k1, d0, d1 = mo.synth_pair(5, 40, 33)
    a = mo.forward(sd, k0, k1, d0, d1, n_layers=2)
    b = mo.forward(sd, k0, k1, d0, d1, n_layers=2, dtype=torch.float64)
    assert float((a["scores"].double() - b["scores"]).abs().max()) < 1e-4


@pytest.mark.skipif(not REF.exists(), reason="/root/reference only exists in the build container")
def test_oracle_against_live_reference_model(pkg):
    sys.path.insert(0, str(REF))
    from lightglue_pytorch_no_plugin import lightglue as ref_lg

    torch.manual_seed(11)
    with torch.no_grad():
        model = ref_lg.LightGlue(features=None, n_layers=2).eval()
        k0, k1, d0, d1 = mo.synth_pair(77, 50, 90)
        r0, r1, rs = model(k0, k1, d0, d1)
        out = mo.forward(model.state_dict(), k0, k1, d0, d1, n_layers=2)
    assert float((out["scores"] - rs).abs().max()) < 5e-5
    assert float((out["desc0"] - r0).abs().max()) < 2e-5 and float((out["desc1"] - r1).abs().max()) < 2e-5
    assert _sha(model.state_dict()) == _sha(pkg.matcher.reference_init_state_dict(11, 2))


def test_weight_repacking_is_exact_in_fp64(pkg):
    """The folded / permuted weights (MatcherWeights) describe the same function: evaluate the packed
    form in FP64 on CPU tensors and compare with the oracle's scores."""
    m = pkg.matcher
    n_layers = 2
    sd = m.reference_init_state_dict(21, n_layers)
    w = m.MatcherWeights({k: v.double() for k, v in sd.items()}, n_layers, device="cpu")
    k0, k1, d0, d1 = mo.synth_pair(9, 30, 41)
    want = mo.forward(sd, k0, k1, d0, d1, n_layers=n_layers, dtype=torch.float64)

    # FP64 re-derivation of the packed weights (MatcherWeights stores FP16; redo the packing in FP64)
    f = {k: v.double() for k, v in sd.items()}
    perm = torch.arange(768).reshape(4, 64, 3).permute(2, 0, 1).reshape(-1)
    c = torch.zeros(256, dtype=torch.float64)
    xs = [d0[0].double(), d1[0].double()]
    encs = [mo.posenc(k.double(), f["posenc.Wr.weight"]) for k in (k0, k1)]

    def heads(t):
        return t.unflatten(-1, (4, 64)).transpose(0, 1)[None]

    def unheads(t):
        return t[0].transpose(0, 1).flatten(-2)

    for i in range(n_layers):
        for kind in ("self_attn", "cross_attn"):
            p = f"transformers.{i}.{kind}."
            if kind == "self_attn":
                w_in, b_in = f[p + "Wqkv.weight"][perm], f[p + "Wqkv.bias"][perm]
                wo, bo = f[p + "out_proj.weight"], f[p + "out_proj.bias"]
            else:
                w_in = torch.cat([f[p + "to_qk.weight"], f[p + "to_v.weight"]], 0)
                b_in = torch.cat([f[p + "to_qk.bias"], f[p + "to_v.bias"]], 0)
                wo, bo = f[p + "to_out.weight"], f[p + "to_out.bias"]
            w1 = f[p + "ffn.0.weight"]
            w1f = torch.cat([w1[:, :256], w1[:, 256:] @ wo], 1)
            b1f = f[p + "ffn.0.bias"] + w1[:, 256:] @ bo + w1[:, :256] @ c
            proj = [x @ w_in.t() + b_in + w_in @ c for x in xs]
            if kind == "self_attn":
                ctx = [unheads(mo.attention(mo.rotary(heads(pj[:, :256]), e), mo.rotary(heads(pj[:, 256:512]), e),
                                            heads(pj[:, 512:]))) for pj, e in zip(proj, encs)]
            else:
                ctx = [unheads(mo.attention(heads(proj[a][:, :256]), heads(proj[1 - a][:, :256]),
                                            heads(proj[1 - a][:, 256:]))) for a in (0, 1)]
            new = []
            for x, cx in zip(xs, ctx):
                hdn = torch.cat([x, cx], -1) @ w1f.t() + b1f
                hdn = torch.nn.functional.gelu(torch.nn.functional.layer_norm(hdn, (512,), f[p + "ffn.1.weight"],
                                                                              f[p + "ffn.1.bias"]))
                new.append(x + hdn @ f[p + "ffn.3.weight"].t())
            xs = new
            c = c + f[p + "ffn.3.bias"]
    np.testing.assert_allclose((xs[0] + c).numpy(), want["desc0"][0].numpy(), atol=1e-10)
    np.testing.assert_allclose((xs[1] + c).numpy(), want["desc1"][0].numpy(), atol=1e-10)
    # and the stored FP16 packing is that FP64 packing, rounded
    assert torch.equal(w.c_final.double(), c.half().double())
    assert w.blocks[0][0].w_in_t.shape == (256, 768) and w.blocks[0][1].w_in_t.shape == (256, 512)
    assert w.wf_t.shape == (256, m.MD_COLS)


def test_flops_per_pair_matches_survey(pkg):
    f = pkg.matcher.flops_per_pair(1024, 1024)
    assert f["attention"] == 576 * 1024 * 1024 * 64  # 36 calls = 576 N M d
    assert abs(f["linear"] / 1e9 - 45.9) < 1.5
