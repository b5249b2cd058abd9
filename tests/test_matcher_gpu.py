"""GPU parity tests of the pair matcher (the caller of the attention path) and its glue kernels,
through the C-ABI of include/lg_match.h, against oracle/matcher_oracle.py and the fixture generated
from the unmodified reference LightGlue (tests/golden/matcher_forward.npz).

Tolerances: the matcher computes in FP16 storage / FP32 accumulate (like the reference's TensorRT
FP16 engine, steps.txt:143), the fixture is the reference's FP32 eager model.  Stated bars:
  * lightglue_scores: max-abs error <= 0.05 log-units and mean-abs <= 0.008 (scores span ~[-20, -5]);
  * lightglue_descriptors: max-abs <= 0.006 relative to the descriptor max-abs;
  * match set (mutual nearest neighbours, threshold 0): >= 95 % of the reference's matches reproduced,
    and every disagreement must be a near-tie in the reference scores (gap < 0.1).
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import matcher_oracle as mo

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

TOL_SCORES_MAX = 0.05
TOL_SCORES_MEAN = 0.008
TOL_DESC_REL = 0.006
MIN_MATCH_AGREEMENT = 0.95


# ------------------------------------------------------------------------------------------------ glue kernels
@pytest.mark.parametrize("rows", [1, 77, 4096])
def test_posenc_kernel(pkg, rows):
    g = torch.Generator().manual_seed(rows)
    kpts = (torch.rand(rows, 2, generator=g) * 2 - 1)
    wr = torch.randn(32, 2, generator=g)
    cs = pkg.matcher.posenc_table(kpts.to(DEV), wr.to(DEV)).cpu()
    proj = kpts.double() @ wr.double().t()
    assert float((cs[..., 0].double() - proj.cos()).abs().max()) < 2e-6
    assert float((cs[..., 1].double() - proj.sin()).abs().max()) < 2e-6


@pytest.mark.parametrize("rows,ld,blocks", [(1, 768, 2), (130, 768, 2), (1000, 512, 1), (257, 1024, 3)])
def test_rotary_kernel_matches_oracle(pkg, rows, ld, blocks):
    g = torch.Generator().manual_seed(rows + ld)
    buf = torch.randn(rows, ld, generator=g).half()
    kpts = torch.rand(1, rows, 2, generator=g) * 2 - 1
    wr = torch.randn(32, 2, generator=g)
    cs = pkg.matcher.posenc_table(kpts[0].to(DEV), wr.to(DEV))
    out = pkg.matcher.rotary_(buf.clone().to(DEV), blocks, cs).cpu()
    enc = mo.posenc(kpts.double(), wr.double())
    for blk in range(blocks):
        t = buf[:, blk * 256:(blk + 1) * 256].double().unflatten(-1, (4, 64)).transpose(0, 1)[None]
        want = mo.rotary(t, enc)[0].transpose(0, 1).flatten(-2)
        got = out[:, blk * 256:(blk + 1) * 256].double()
        assert float((got - want).abs().max()) <= 2e-3  # FP16 rounding of values up to ~6
        assert float((got - want.half().double()).abs().max()) <= 4e-3
    assert torch.equal(out[:, blocks * 256:], buf[:, blocks * 256:])  # the rest of the row is untouched


@pytest.mark.parametrize("rows,ld", [(1, 512), (9, 512), (1031, 512), (64, 640)])
def test_layernorm_gelu_kernel(pkg, rows, ld):
    g = torch.Generator().manual_seed(rows)
    h = (torch.randn(rows, ld, generator=g) * 3 + 0.5).half()
    gamma = torch.randn(512, generator=g)
    beta = torch.randn(512, generator=g)
    hd = h.clone().to(DEV)
    pkg.matcher.layernorm_gelu_(hd[:, :512], gamma.to(DEV), beta.to(DEV))
    want = F.gelu(F.layer_norm(h[:, :512].double(), (512,), gamma.double(), beta.double(), 1e-5))
    got = hd.cpu()
    assert float((got[:, :512].double() - want).abs().max()) <= 4e-3
    assert torch.equal(got[:, 512:], h[:, 512:])


@pytest.mark.parametrize("rows,n_blocks,n_rot,ldx,ldo", [(1, 3, 2, 256, 768), (77, 3, 2, 512, 768), (1031, 2, 0, 256, 512), (2048, 3, 2, 512, 768),
                                                         (20000, 3, 2, 512, 1024), (20001, 2, 0, 512, 768), (9000, 1, 1, 256, 256)])
def test_fused_projection_rotary_kernel(pkg, rows, n_blocks, n_rot, ldx, ldo):
    """Wqkv / to_qk|to_v projection with the rotary in the GEMM epilogue (lightglue.py:139-147, 190-193) against an FP64 restatement
    (x @ W^T + b, then oracle rotary on the first n_rot blocks): ragged last tile, several items per persistent CTA, strided rows
    with a sentinel in the gaps and past the last row."""
    g = torch.Generator().manual_seed(rows + n_blocks)
    n = 256 * n_blocks
    x = torch.randn(rows, ldx, generator=g).half()
    w = (torch.randn(n, 256, generator=g) / 16 * 1.5).half()
    bias = (torch.randn(n, generator=g) * 0.5).half()
    kpts = torch.rand(1, rows, 2, generator=g) * 2 - 1
    wr = torch.randn(32, 2, generator=g)
    cs = pkg.matcher.posenc_table(kpts[0].to(DEV), wr.to(DEV))
    guard = 3
    out = torch.full((rows + guard, ldo), -7.0).half()
    od = out.clone().to(DEV)
    pkg.matcher.proj_rotary(x.to(DEV)[:, :256], w.to(DEV), bias.to(DEV), n_rot, cs if n_rot else None, od[:rows, :n])
    got = od.cpu()
    want = x[:, :256].double() @ w.double().t() + bias.double()
    enc = mo.posenc(kpts.double(), wr.double())
    for blk in range(n_rot):
        t = want[:, blk * 256:(blk + 1) * 256].unflatten(-1, (4, 64)).transpose(0, 1)[None]
        want[:, blk * 256:(blk + 1) * 256] = mo.rotary(t, enc)[0].transpose(0, 1).flatten(-2)
    err = (got[:rows, :n].double() - want).abs()
    assert bool((err <= 6e-4 * want.abs() + 3e-4).all()), float(err.max())   # FP16 rounding of the result (2^-11 relative) + cos / sin in FP32
    assert torch.equal(got[rows:], out[rows:]) and torch.equal(got[:, n:], out[:, n:])   # nothing written outside [rows, n]


@pytest.mark.parametrize("rows,ldx,ldh", [(1, 512, 512), (77, 512, 512), (128, 512, 512), (1031, 512, 640), (4736, 512, 512), (4737, 768, 512), (20000, 512, 512), (40000, 768, 512)])
def test_fused_ffn1_ln_gelu_kernel(pkg, rows, ldx, ldh):
    """ffn[0..2] (Linear 512->512, LayerNorm, erf GELU; lightglue.py:131-136) as one tcgen05 kernel against an FP64 restatement:
    ragged last tile, several tiles per persistent CTA (20000 rows = 157 tiles on 148 SMs), strided input / output rows with a
    sentinel in the gaps and past the last row.  Up to 37 tiles (4736 rows) the launcher takes the few-rows kernel (4-CTA clusters
    splitting the 512 columns, statistics exchanged through distributed shared memory), above that the persistent one."""
    g = torch.Generator().manual_seed(rows + ldx)
    x = torch.randn(rows, ldx, generator=g).half()
    w = (torch.randn(512, 512, generator=g) / 512 ** 0.5 * 1.5).half()
    bias = (torch.randn(512, generator=g) * 0.5).half()
    gamma = torch.randn(512, generator=g)
    beta = torch.randn(512, generator=g)
    guard = 3
    out = torch.full((rows + guard, ldh), -7.0).half()
    od = out.clone().to(DEV)
    pkg.matcher.ffn1_ln_gelu(x.to(DEV)[:, :512], w.to(DEV), bias.to(DEV), gamma.to(DEV), beta.to(DEV), od[:rows, :512])
    got = od.cpu()
    pre = x[:, :512].double() @ w.double().t() + bias.double()
    want = F.gelu(F.layer_norm(pre, (512,), gamma.double(), beta.double(), 1e-5))
    # Like the library GEMM + lg_match_layernorm_gelu path, the kernel holds the pre-LayerNorm values in FP16 (statistics are taken
    # in FP32 before that rounding).  Bars: against the same computation on FP16-rounded pre-activations only the output rounding
    # (2^-11 relative) and the erf approximation remain; against exact arithmetic the rounding of values up to ~6 adds up to
    # ~5e-3 after the affine (|gamma| up to ~3.5).
    want_h = F.gelu(F.layer_norm(pre.half().double(), (512,), gamma.double(), beta.double(), 1e-5))
    g = got[:rows, :512].double()
    # (elementwise, except where the FP32 accumulator and the FP64 product round to different FP16 neighbours: rare, <= 1 ulp of
    # the pre-activation)
    off = ((g - want_h).abs() > 6e-4 * want_h.abs() + 4e-4).double().mean()
    assert float(off) <= 2e-3, float(off)
    err = (g - want).abs()
    assert float(err.max()) <= 1.2e-2, float(err.max())
    assert float(err.mean()) <= 6e-4, float(err.mean())
    assert torch.equal(got[rows:], out[rows:]) and torch.equal(got[:, 512:], out[:, 512:])   # nothing written outside [rows, 512]


@pytest.mark.parametrize("rows", [300, 20000, 40001])
def test_fused_ffn1_cta_pair_variant_is_bit_equal(pkg, rows):
    """The persistent ffn-up kernel as CTA pairs (cta_group::2: M = 256 MMAs, half of every W1' tile per CTA, TMA bytes of both
    CTAs completing on the leader's barrier, multicast commits) computes exactly what the one-CTA variant computes."""
    g = torch.Generator().manual_seed(rows)
    x = torch.randn(rows, 512, generator=g).half().to(DEV)
    w = (torch.randn(512, 512, generator=g) / 512 ** 0.5 * 1.5).half().to(DEV)
    bias = (torch.randn(512, generator=g) * 0.5).half().to(DEV)
    gamma, beta = torch.randn(512, generator=g).to(DEV), torch.randn(512, generator=g).to(DEV)
    lib = pkg.load()
    # rows <= 4736 would take the few-rows cluster kernel: force the persistent one by going through a strided view?  No — the
    # launcher picks by tile count only, so 300 rows exercises the few-rows kernel in both settings (must still agree).
    outs = []
    try:
        for pair in (0, 1):
            lib.lg_match_debug_set_ffn_pair(pair)
            o = torch.full((rows + 2, 512), -3.0, dtype=torch.float16, device=DEV)
            pkg.matcher.ffn1_ln_gelu(x, w, bias, gamma, beta, o[:rows])
            torch.cuda.synchronize()
            outs.append(o.cpu())
    finally:
        lib.lg_match_debug_set_ffn_pair(-1)   # back to the choice by size
    assert torch.equal(outs[0], outs[1])
    assert bool((outs[1][rows:] == -3.0).all())


@pytest.mark.parametrize("b,n0,n1", [(1, 1, 1), (1, 1, 70), (1, 160, 130), (3, 257, 64), (2, 512, 515), (1, 1024, 1024)])
def test_assignment_kernels_match_oracle(pkg, b, n0, n1):
    g = torch.Generator().manual_seed(n0 * 3 + n1)
    sim = torch.randn(b, n0, n1, generator=g) * 3
    z = torch.randn(b * (n0 + n1), 8, generator=g).half()  # logits live in a strided column, like MD[:, 256]
    z0 = z[: b * n0, 3].unflatten(0, (b, n0))
    z1 = z[b * n0:, 3].unflatten(0, (b, n1))
    zd = z.to(DEV)
    m0, ms, scores = pkg.matcher.assign(sim.to(DEV), zd[: b * n0, 3].unflatten(0, (b, n0)),
                                        zd[b * n0:, 3].unflatten(0, (b, n1)), 0.0, want_scores=True)
    cert = F.logsigmoid(z0.double())[:, :, None] + F.logsigmoid(z1.double())[:, None, :]
    want = F.log_softmax(sim.double(), 2) + F.log_softmax(sim.double(), 1) + cert
    assert float((scores.cpu().double() - want).abs().max()) < 2e-5
    wm0, wms = mo.filter_matches(want, 0.0)
    assert torch.equal(m0.cpu().long(), wm0)
    np.testing.assert_allclose(ms.cpu().numpy(), wms.float().numpy(), rtol=2e-5, atol=1e-30)
    # with a threshold: identical to thresholding the oracle's mscores
    thr = float(wms[wms > 0].median()) if (wms > 0).any() else 0.5
    m0t, _, _ = pkg.matcher.assign(sim.to(DEV), zd[: b * n0, 3].unflatten(0, (b, n0)),
                                   zd[b * n0:, 3].unflatten(0, (b, n1)), thr)
    wm0t, _ = mo.filter_matches(want.float(), thr)
    assert int((m0t.cpu().long() != wm0t).sum()) <= 1  # a score exactly at the threshold may flip


# ------------------------------------------------------------------------------------------------ full matcher
def _agreement(ref_scores: torch.Tensor, got_m0: torch.Tensor):
    """share of the reference's mutual-NN matches reproduced + the largest reference margin among disagreements.  A reference
    match (i, j) needs j = argmax of row i AND i = argmax of column j; it can only be lost to score noise if one of those two
    was nearly tied in the reference, so the margin of a disagreement is min(row margin, column margin) to the runner-up: a
    disagreement is benign when that is small (scores agree to TOL_SCORES_MAX), a large one is a real error."""
    want_m0, _ = mo.filter_matches(ref_scores, 0.0)
    want, got = want_m0[0], got_m0[0].long().cpu()
    has = want >= 0
    agree = float((got[has] == want[has]).float().mean()) if has.any() else 1.0
    worst_gap = 0.0
    sc = ref_scores[0]
    for i in torch.nonzero(has & (got != want))[:, 0].tolist():
        j = int(want[i])
        row, col = sc[i], sc[:, j]
        row_margin = float(row[j] - row.topk(2).values[1]) if row.numel() > 1 else float("inf")
        col_margin = float(col[i] - col.topk(2).values[1]) if col.numel() > 1 else float("inf")
        worst_gap = max(worst_gap, min(row_margin, col_margin))
    return agree, worst_gap


def test_matcher_reproduces_reference_fixture(pkg, golden_dir):
    z = np.load(golden_dir / "matcher_forward.npz")
    for name in sorted({k.split(".")[0] for k in z.files}):
        wseed, n_layers, fseed, n0, n1 = (int(v) for v in z[f"{name}.meta"])
        matcher = pkg.PairMatcher(n_layers=n_layers, seed=wseed, threshold=0.0, device=DEV)
        k0, k1, d0, d1 = (t.to(DEV) for t in mo.synth_pair(fseed, n0, n1))
        out = matcher.match(k0, k1, d0, d1, want_scores=True)
        ref_scores = torch.from_numpy(z[f"{name}.scores"])
        err = (out["scores"].cpu() - ref_scores).abs()
        dmax = float(np.abs(z[f"{name}.desc0"]).max())
        e0 = float((out["desc0"].cpu() - torch.from_numpy(z[f"{name}.desc0"])).abs().max()) / dmax
        e1 = float((out["desc1"].cpu() - torch.from_numpy(z[f"{name}.desc1"])).abs().max()) / dmax
        agree, gap = _agreement(ref_scores, out["matches0"])
        print(f"{name}: scores max {float(err.max()):.4f} mean {float(err.mean()):.5f}; desc rel {e0:.4f} {e1:.4f}; "
              f"match agreement {agree:.3f} worst gap {gap:.4f}")
        assert float(err.max()) <= TOL_SCORES_MAX and float(err.mean()) <= TOL_SCORES_MEAN
        assert e0 <= TOL_DESC_REL and e1 <= TOL_DESC_REL
        assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1


def test_batched_ragged_pairs_match_oracle(pkg):
    """B pairs at once, N0 != N1, non-tile-multiple: every pair equals the FP32 oracle on that pair,
    and equals the same pair matched alone (bit-exact: pairs are independent)."""
    n_layers, b, n0, n1 = 3, 3, 200, 333
    sd = pkg.reference_init_state_dict(4, n_layers)
    matcher = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV)
    feats = [mo.synth_pair(50 + i, n0, n1) for i in range(b)]
    k0, k1, d0, d1 = (torch.cat([f[j] for f in feats]).to(DEV) for j in range(4))
    out = matcher.match(k0, k1, d0, d1, want_scores=True)
    scores, m0 = out["scores"].cpu().clone(), out["matches0"].cpu().clone()
    for i, f in enumerate(feats):
        want = mo.forward(sd, *f, n_layers=n_layers, threshold=0.0)
        err = (scores[i] - want["scores"][0]).abs()
        assert float(err.max()) <= TOL_SCORES_MAX and float(err.mean()) <= TOL_SCORES_MEAN
        agree, gap = _agreement(want["scores"], m0[i:i + 1])
        assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1
        alone = matcher.match(*(t.to(DEV) for t in f), want_scores=True)
        assert torch.equal(alone["scores"].cpu()[0], scores[i])
        assert torch.equal(alone["matches0"].cpu()[0], m0[i])


def test_fused_and_unfused_ffn_paths_agree(pkg, monkeypatch):
    """The same forward through the fused ffn-up kernel and through library GEMM + lg_match_layernorm_gelu (LG_MATCH_FUSED_FFN=0),
    at a shape big enough for the fused path to be chosen: scores agree to FP16-path noise, match sets almost entirely."""
    n_layers, b, n = 3, 6, 1100   # 13 200 rows = 104 tiles: fused by the heuristic
    sd = pkg.reference_init_state_dict(5, n_layers)
    feats = [mo.synth_pair(300 + i, n, n) for i in range(b)]
    f = [torch.cat([x[j] for x in feats]).to(DEV) for j in range(4)]
    assert pkg.load().lg_match_ffn1_prefer_fused(2 * b * n) == 1
    a = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV)
    assert a.fused_ffn
    ra = a.match(*f, want_scores=True)
    sa, ma, la = ra["scores"].clone(), ra["matches0"].clone(), ra["launches"]
    monkeypatch.setenv("LG_MATCH_FUSED_FFN", "0")
    c = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV)
    assert not c.fused_ffn
    rc = c.match(*f, want_scores=True)
    assert rc["launches"] == la + 2 * n_layers
    assert float((rc["scores"] - sa).abs().max()) <= 0.03
    assert float((rc["matches0"] == ma).float().mean()) >= 0.98
    # and against the FP32 CPU oracle, first pair
    want = mo.forward(sd, *feats[0], n_layers=n_layers, threshold=0.0)
    assert float((sa[0].cpu() - want["scores"][0]).abs().max()) <= TOL_SCORES_MAX


def test_fused_and_unfused_projection_paths_agree(pkg, monkeypatch):
    """The same forward with the input projections through lg_match_proj_rotary (rotary in the GEMM epilogue) and through library
    GEMM + lg_match_rotary (LG_MATCH_FUSED_PROJ=0): scores agree to FP16-path noise, and both agree with the FP32 CPU oracle."""
    n_layers, b, n0, n1 = 3, 7, 1100, 1000   # 14 700 rows: 115 row tiles (fused by the heuristic), ragged, both cross directions as separate launches
    sd = pkg.reference_init_state_dict(6, n_layers)
    feats = [mo.synth_pair(400 + i, n0, n1) for i in range(b)]
    f = [torch.cat([x[j] for x in feats]).to(DEV) for j in range(4)]
    a = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV)
    ra = a.match(*f, want_scores=True)
    assert a.plan(b, n0, n1, True).fused_proj_self and a.plan(b, n0, n1, True).fused_proj_cross
    sa, ma, la = ra["scores"].clone(), ra["matches0"].clone(), ra["launches"]
    monkeypatch.setenv("LG_MATCH_FUSED_PROJ", "0")
    c = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV)
    rc = c.match(*f, want_scores=True)
    assert not c.plan(b, n0, n1, True).fused_proj_self
    assert rc["launches"] == la + n_layers
    assert float((rc["scores"] - sa).abs().max()) <= 0.03
    assert float((rc["matches0"] == ma).float().mean()) >= 0.98
    want = mo.forward(sd, *feats[0], n_layers=n_layers, threshold=0.0)
    assert float((sa[0].cpu() - want["scores"][0]).abs().max()) <= TOL_SCORES_MAX


def test_graph_replay_equals_eager(pkg):
    n_layers = 2
    sd = pkg.reference_init_state_dict(8, n_layers)
    f = [t.to(DEV) for t in mo.synth_pair(123, 300, 280)]
    a = pkg.PairMatcher(sd, n_layers=n_layers, device=DEV, use_graph=True).match(*f, want_scores=True)
    b = pkg.PairMatcher(sd, n_layers=n_layers, device=DEV, use_graph=False).match(*f, want_scores=True)
    assert torch.equal(a["scores"], b["scores"]) and torch.equal(a["matches0"], b["matches0"])
    assert a["launches"] == b["launches"] > 0


def test_full_size_pair_properties(pkg):
    """BASELINE configs[4] size (2048 keypoints), too slow for the CPU oracle: size-independent
    properties.  (1) matches0 / mscores0 are consistent with the returned score matrix (mutual argmax);
    (2) swapping the two images transposes the score matrix; (3) batch order is irrelevant."""
    n_layers, b, n = 9, 2, 2048
    matcher = pkg.PairMatcher(n_layers=n_layers, seed=0, threshold=0.0, device=DEV)
    feats = [mo.synth_pair(900 + i, n, n, width=1280, height=960) for i in range(b)]
    k0, k1, d0, d1 = (torch.cat([f[j] for f in feats]).to(DEV) for j in range(4))
    out = matcher.match(k0, k1, d0, d1, want_scores=True)
    scores, m0, ms = out["scores"].clone(), out["matches0"].clone(), out["mscores0"].clone()
    assert torch.isfinite(scores).all()
    wm0, wms = mo.filter_matches(scores, 0.0)
    assert torch.equal(m0.long(), wm0)
    torch.testing.assert_close(ms, wms, rtol=1e-5, atol=0)
    swapped = matcher.match(k1, k0, d1, d0, want_scores=True)["scores"]
    assert float((swapped.transpose(1, 2) - scores).abs().max()) <= 2e-2  # different GEMM/attention row order
    flipped = matcher.match(k0.flip(0), k1.flip(0), d0.flip(0), d1.flip(0), want_scores=True)
    assert torch.equal(flipped["scores"].flip(0), scores) and torch.equal(flipped["matches0"].flip(0), m0)


def test_matcher_refuses_cpu_tensors(pkg):
    matcher = pkg.PairMatcher(n_layers=1, device=DEV)
    f = mo.synth_pair(1, 8, 8)
    with pytest.raises(RuntimeError):
        matcher.match(*f)


def test_nine_layer_ragged_pair_matches_oracle(pkg):
    """Full depth (9 layers), ragged and non-tile-multiple keypoint counts, against the FP32 CPU oracle."""
    sd = pkg.reference_init_state_dict(0, 9)
    f = mo.synth_pair(31337, 600, 517)
    out = pkg.PairMatcher(sd, n_layers=9, threshold=0.0, device=DEV).match(*(t.to(DEV) for t in f), want_scores=True)
    want = mo.forward(sd, *f, n_layers=9, threshold=0.0)
    err = (out["scores"].cpu() - want["scores"]).abs()
    assert float(err.max()) <= TOL_SCORES_MAX and float(err.mean()) <= TOL_SCORES_MEAN
    agree, gap = _agreement(want["scores"], out["matches0"])
    assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1


@pytest.mark.parametrize("n", [1024, 2048])
def test_config4_full_size_pair_matches_oracle(pkg, n, capsys):
    """BASELINE.json configs[3] (the 1x1024x1024 engine shape) and the 2048-keypoint shape of configs[4]: the full 9-layer
    matcher against the FP32 CPU oracle (oracle/matcher_oracle.py = lightglue_pytorch_no_plugin/lightglue.py:359-384 +
    filter_matches :265-288) on freiburg-shaped synthetic SuperPoint features.  Prints the agreement figures DESIGN.md quotes."""
    sd = pkg.reference_init_state_dict(0, 9)
    f = mo.synth_pair(1000 + n, n, n) if n == 1024 else mo.synth_pair(1000 + n, n, n, width=1280, height=960)
    out = pkg.PairMatcher(sd, n_layers=9, threshold=0.0, device=DEV).match(*(t.to(DEV) for t in f), want_scores=True)
    want = mo.forward(sd, *f, n_layers=9, threshold=0.0)
    err = (out["scores"].cpu() - want["scores"]).abs()
    agree, gap = _agreement(want["scores"], out["matches0"])
    n_ref = int((want["matches0"] >= 0).sum())
    with capsys.disabled():
        print(f"\n[config4 parity] {n}x{n} kpts, 9 layers: scores max-abs err {float(err.max()):.4f}, mean-abs {float(err.mean()):.5f}; "
              f"mutual-NN matches in the reference {n_ref}, reproduced {agree * 100:.2f} % (largest reference score gap among "
              f"disagreements {gap:.4f})")
    assert float(err.max()) <= TOL_SCORES_MAX and float(err.mean()) <= TOL_SCORES_MEAN
    assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1


def test_batched_full_depth_pairs_through_the_fused_ffn_match_oracle(pkg, capsys):
    """The batched configuration the pairs/s numbers are quoted on runs the fused ffn-up kernel (a single pair does not: too few
    rows, lg_match_ffn1_prefer_fused): 8 pairs x 1024 keypoints, all 9 layers, every pair against the FP32 CPU oracle."""
    b, n = 8, 1024
    sd = pkg.reference_init_state_dict(0, 9)
    feats = [mo.synth_pair(2000 + i, n, n) for i in range(b)]
    matcher = pkg.PairMatcher(sd, n_layers=9, threshold=0.0, device=DEV)
    out = matcher.match(*(torch.cat([f[j] for f in feats]).to(DEV) for j in range(4)), want_scores=True)
    assert matcher.fused_ffn and matcher.plan(b, n, n, True).prefer_fused_ffn
    worst_max = worst_mean = 0.0
    n_ref = n_same = 0
    for i, f in enumerate(feats):
        want = mo.forward(sd, *f, n_layers=9, threshold=0.0)
        err = (out["scores"][i:i + 1].cpu() - want["scores"]).abs()
        worst_max, worst_mean = max(worst_max, float(err.max())), max(worst_mean, float(err.mean()))
        agree, gap = _agreement(want["scores"], out["matches0"][i:i + 1])
        assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1, (i, agree, gap)
        ref_m = want["matches0"][0]
        n_ref += int((ref_m >= 0).sum())
        n_same += int(((out["matches0"][i].cpu().long() == ref_m) & (ref_m >= 0)).sum())
    with capsys.disabled():
        print(f"\n[fused ffn parity] {b} pairs x {n} kpts, 9 layers: scores max-abs err {worst_max:.4f}, worst mean-abs {worst_mean:.5f}; "
              f"mutual-NN matches in the reference {n_ref}, reproduced {n_same} ({100.0 * n_same / max(n_ref, 1):.2f} %)")
    assert worst_max <= TOL_SCORES_MAX and worst_mean <= TOL_SCORES_MEAN


def test_assignment_survives_nan_and_inf_rows(pkg):
    """FP16 overflow upstream can hand the assignment a similarity row of NaN / -inf: such a row has no argmax (the index stays
    at the INT_MAX sentinel), must not be dereferenced (illegal address) and must report 'no match'; other rows are unaffected."""
    g = torch.Generator().manual_seed(5)
    b, n0, n1 = 2, 130, 96
    sim = torch.randn(b, n0, n1, generator=g) * 3
    z0, z1 = torch.randn(b, n0, generator=g).half(), torch.randn(b, n1, generator=g).half()
    clean = pkg.matcher.assign(sim.to(DEV), z0.to(DEV), z1.to(DEV), 0.0)
    bad = sim.clone()
    bad[0, 7] = float("nan")
    bad[1, 100] = float("-inf")
    m0, ms, _ = pkg.matcher.assign(bad.to(DEV), z0.to(DEV), z1.to(DEV), 0.0)
    torch.cuda.synchronize()
    assert int(m0[0, 7]) == -1 and int(m0[1, 100]) == -1
    assert float(ms[1, 100]) == 0.0
    assert ((m0.cpu() >= -1) & (m0.cpu() < n1)).all()
    # a -inf row adds exp(-inf) = 0 to every column's log-sum-exp and can win no column: the other rows of that pair must come
    # out exactly as if the row did not exist
    keep = torch.ones(n0, dtype=torch.bool)
    keep[100] = False
    w0, ws, _ = pkg.matcher.assign(bad[1:2, keep].contiguous().to(DEV), z0[1:2, keep].contiguous().to(DEV), z1[1:2].to(DEV), 0.0)
    assert torch.equal(m0[1].cpu()[keep], w0[0].cpu())
    torch.testing.assert_close(ms[1].cpu()[keep], ws[0].cpu(), rtol=1e-5, atol=1e-30)
    assert torch.isfinite(ms[1]).all()
    del clean


def test_plan_caches_are_bounded_and_alternating_shapes_do_not_replan(pkg):
    """The engine is dynamic (every pair has its own keypoint counts): shape plans are cached LRU with a bound, results do not
    depend on what was evicted, and two alternating shapes are swapped instead of rebuilt (native engine: plans_resident)."""
    sd = pkg.reference_init_state_dict(1, 2)
    pm = pkg.PairMatcher(sd, n_layers=2, threshold=0.0, device=DEV, max_plans=3)
    shapes = [(40, 33), (64, 64), (50, 70), (33, 40), (90, 10)]
    feats = {s: tuple(t.to(DEV) for t in mo.synth_pair(7 + i, *s)) for i, s in enumerate(shapes)}
    first = {s: pm.match(*feats[s])["matches0"].clone() for s in shapes}
    assert len(pm._plans) == 3
    for s in shapes:                     # every shape again, most of them after an eviction
        assert torch.equal(pm.match(*feats[s])["matches0"], first[s])
    assert len(pm._plans) == 3
    # native engine: active + parked plans, LRU eviction at 4, alternation without re-planning
    eng = pkg.NativeMatcher(sd, n_layers=2, device=0, threshold=0.0)
    lib = pkg.load()
    host = {s: mo.synth_pair(7 + i, *s) for i, s in enumerate(shapes)}
    got = {}
    for s in shapes:
        k0, k1, d0, d1 = host[s]
        got[s] = eng.match_host(k0.numpy(), k1.numpy(), d0.half().numpy(), d1.half().numpy())["matches0"].copy()
        assert float((got[s] == first[s].cpu().numpy()).mean()) >= MIN_MATCH_AGREEMENT
    assert lib.lg_matcher_plans_resident(eng._h) == 4
    for _ in range(3):
        for s in shapes[-2:]:
            k0, k1, d0, d1 = host[s]
            assert np.array_equal(eng.match_host(k0.numpy(), k1.numpy(), d0.half().numpy(), d1.half().numpy())["matches0"], got[s])
    assert lib.lg_matcher_plans_resident(eng._h) == 4


# ------------------------------------------------------------------------------------------------ native C++ engine
def test_native_engine_reproduces_reference_fixture(pkg, golden_dir):
    """lg_matcher_* (C++ / cuBLASLt, no torch in the call path) against the reference-generated fixture: the
    match set must agree with the reference's, and with the Python-driven matcher on the same weights."""
    z = np.load(golden_dir / "matcher_forward.npz")
    for name in sorted({k.split(".")[0] for k in z.files}):
        wseed, n_layers, fseed, n0, n1 = (int(v) for v in z[f"{name}.meta"])
        sd = pkg.reference_init_state_dict(wseed, n_layers)
        eng = pkg.NativeMatcher(sd, n_layers=n_layers, threshold=0.0, device=0)
        k0, k1, d0, d1 = mo.synth_pair(fseed, n0, n1)
        out = eng.match(k0.to(DEV), k1.to(DEV), d0.half().to(DEV), d1.half().to(DEV))
        ref_scores = torch.from_numpy(z[f"{name}.scores"])
        agree, gap = _agreement(ref_scores, out["matches0"])
        assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1, (name, agree, gap)
        # mscores0 of the agreed matches: exp(score) within the score tolerance
        want_m0, want_ms = mo.filter_matches(ref_scores, 0.0)
        same = (out["matches0"].cpu().long() == want_m0) & (want_m0 >= 0)
        rel = (out["mscores0"].cpu()[same] / want_ms[same] - 1).abs()
        assert rel.numel() == 0 or float(rel.max()) <= 2 * TOL_SCORES_MAX
        py = pkg.PairMatcher(sd, n_layers=n_layers, threshold=0.0, device=DEV).match(
            k0.to(DEV), k1.to(DEV), d0.to(DEV), d1.to(DEV))
        assert float((py["matches0"] == out["matches0"]).float().mean()) >= 0.98
        eng.close()


def test_native_engine_host_buffers_batches_and_replanning(pkg):
    n_layers = 2
    sd = pkg.reference_init_state_dict(9, n_layers)
    eng = pkg.NativeMatcher(sd, n_layers=n_layers, threshold=0.0, device=0)
    for (b, n0, n1) in [(3, 200, 333), (2, 256, 256), (1, 1, 70)]:
        feats = [mo.synth_pair(70 + i, n0, n1) for i in range(b)]
        k0, k1, d0, d1 = (torch.cat([f[j] for f in feats]) for j in range(4))
        launches = eng.plan(b, n0, n1)
        # per layer: 2 x (in-proj, [rotary,] attention x 1 or 2, ffn-up (fused: 1 launch, else GEMM + LayerNorm/GELU), ffn-down)
        fused = pkg.load().lg_match_ffn1_prefer_fused(b * (n0 + n1))
        fused_proj = pkg.load().lg_match_proj_prefer_fused(b * (n0 + n1), 3, 2)   # Wqkv + rotary as one launch
        assert launches == (1 + n_layers * ((13 if n0 != n1 else 11) - (2 if fused else 0) - (1 if fused_proj else 0)) + 9)
        host = eng.match_host(k0.numpy(), k1.numpy(), d0.half().numpy(), d1.half().numpy())
        dev = eng.match(k0.to(DEV), k1.to(DEV), d0.half().to(DEV), d1.half().to(DEV))
        torch.cuda.synchronize()
        assert np.array_equal(host["matches0"], dev["matches0"].cpu().numpy())
        assert np.array_equal(host["mscores0"], dev["mscores0"].cpu().numpy())
        for i, f in enumerate(feats):
            want = mo.forward(sd, *f, n_layers=n_layers, threshold=0.0)
            agree, gap = _agreement(want["scores"], dev["matches0"][i:i + 1])
            assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1
    eng.close()


def test_native_engine_async_submit_pipeline_matches_synchronous_path(pkg):
    """lg_matcher_submit_host / lg_matcher_sync: several sub-batches in flight (two staging slots reused, a ragged
    last sub-batch that re-plans) give exactly what the synchronous host call gives for each sub-batch."""
    n_layers = 2
    sd = pkg.reference_init_state_dict(4, n_layers)
    eng = pkg.NativeMatcher(sd, n_layers=n_layers, threshold=0.0, device=0)
    ref = pkg.NativeMatcher(sd, n_layers=n_layers, threshold=0.0, device=0)
    pairs, sub, n = 11, 3, 192
    feats = [mo.synth_pair(300 + i, n, n) for i in range(pairs)]
    k0, k1, d0, d1 = (torch.cat([f[j] for f in feats]) for j in range(4))
    k0, k1 = k0.contiguous().pin_memory(), k1.contiguous().pin_memory()
    d0, d1 = d0.half().contiguous().pin_memory(), d1.half().contiguous().pin_memory()
    m0 = torch.full((pairs, n), -7, dtype=torch.int32).pin_memory()
    ms = torch.full((pairs, n), -7.0, dtype=torch.float32).pin_memory()
    for _ in range(2):                      # second round reuses the slots after a drain
        m0.fill_(-7)
        for a in range(0, pairs, sub):
            b = min(pairs, a + sub)
            eng.submit_host(k0[a:b], k1[a:b], d0[a:b], d1[a:b], m0[a:b], ms[a:b])
        eng.sync()
        for a in range(0, pairs, sub):
            b = min(pairs, a + sub)
            want = ref.match_host(k0[a:b].numpy(), k1[a:b].numpy(), d0[a:b].numpy(), d1[a:b].numpy())
            assert np.array_equal(m0[a:b].numpy(), want["matches0"])
            assert np.array_equal(ms[a:b].numpy(), want["mscores0"])
    eng.close()
    ref.close()


def test_native_engine_survives_a_failed_plan(pkg):
    """A plan that cannot be allocated returns an error code (no abort, no half-built plan); the engine then
    plans and matches a normal shape as if nothing had happened."""
    n_layers = 1
    sd = pkg.reference_init_state_dict(2, n_layers)
    eng = pkg.NativeMatcher(sd, n_layers=n_layers, threshold=0.0, device=0)
    with pytest.raises(RuntimeError):
        eng.plan(4096, 8192, 8192)          # the score matrix alone would be 1 TiB
    assert pkg.load().lg_matcher_launches(eng._h) == 0
    k0, k1, d0, d1 = mo.synth_pair(5, 96, 80)
    got = eng.match_host(k0.numpy(), k1.numpy(), d0.half().numpy(), d1.half().numpy())
    want = mo.forward(sd, k0, k1, d0, d1, n_layers=n_layers, threshold=0.0)
    agree, gap = _agreement(want["scores"], torch.from_numpy(got["matches0"]))
    assert agree >= MIN_MATCH_AGREEMENT and gap < 0.1
    eng.close()


def test_native_pair_bench_binary(pkg, tmp_path):
    """The C++ harness (csrc/tools/lg_pair_bench.cpp) end to end: blob file -> forked worker -> JSON line."""
    import json
    import subprocess

    exe = pkg._binding.lib_path().parent / "lg_pair_bench"
    assert exe.exists(), "build() did not produce lg_pair_bench"
    m = pkg.matcher
    blob = tmp_path / "w.blob"
    blob.write_bytes(m.weights_blob(m.MatcherWeights(m.reference_init_state_dict(0, 2), 2, device="cpu")))
    r = subprocess.run([str(exe), "--blob", str(blob), "--layers", "2", "--gpus", "1", "--kpts", "256", "--pairs", "6",
                        "--sub", "4", "--steps", "2"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["unit"] == "pairs/s" and line["value"] > 0 and line["n_gpus"] == 1 and line["kpts"] == 256
    assert 0 < line["mutual_nn_per_pair"] <= 256
