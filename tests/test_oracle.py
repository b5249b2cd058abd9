"""CPU: the oracle against the reference-generated golden fixtures (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import attention_oracle as orc


def _cases(npz):
    tags = sorted({k.rsplit("_", 1)[0] for k in npz.files if k.endswith("_q")})
    assert tags
    return tags


def test_oracle_matches_reference_attention_forward(golden_dir):
    """oracle == lightglue_pytorch_no_plugin Attention.forward (lightglue.py:75-85) outputs."""
    g = np.load(golden_dir / "attn_forward.npz")
    for tag in _cases(g):
        q, k, v, o = (g[f"{tag}_{x}"] for x in "qkvo")
        got64 = orc.attention_np(q, k, v, np.float64)
        got32 = orc.attention_np(q, k, v, np.float32)
        # reference ran in FP32 on CPU: agreement to FP32 rounding of a <=257-term softmax sum
        assert orc.max_abs_err(got64, o) < 2e-6, tag
        assert orc.max_abs_err(got32, o) < 4e-6, tag
        assert got64.shape == o.shape


def test_oracle_matches_reference_plugin_twin(golden_dir):
    """oracle == MHAHeadDim64.forward (eager SDPA twin of the plugin, with_plugin/lightglue.py:35-46)."""
    g = np.load(golden_dir / "attn_forward.npz")
    s = np.load(golden_dir / "attn_sdpa.npz")
    for tag in _cases(g):
        q, k, v = (g[f"{tag}_{x}"] for x in "qkv")
        assert orc.max_abs_err(orc.attention_np(q, k, v), s[f"{tag}_o"]) < 2e-6, tag


def test_oracle_matches_reference_callsites(golden_dir):
    """oracle == what the reference model computed at its own attention call sites (FP32 inputs)."""
    g = np.load(golden_dir / "callsites.npz")
    tags = _cases(g)
    assert len(tags) == 8  # 2 layers x (self0, self1, cross01, cross10)
    for tag in tags:
        q, k, v, o = (g[f"{tag}_{x}"] for x in "qkvo")
        assert q.shape[1:] == (4, q.shape[2], 64)
        err = orc.max_abs_err(orc.attention_np(q, k, v), o)
        assert err < 1e-5 * max(1.0, float(np.abs(o).max())), (tag, err)
    # cross-attention is ragged: N != M (64 vs 48 keypoints)
    assert g["L0_cross01_q"].shape[2] == 64 and g["L0_cross01_k"].shape[2] == 48
    assert g["L0_cross10_q"].shape[2] == 48 and g["L0_cross10_k"].shape[2] == 64


def test_oracle_torch_port_equals_numpy():
    import torch

    q, k, v = orc.make_qkv(3, 1, 4, 70, 33)
    t = orc.attention_torch_cpu(*(torch.from_numpy(x.astype(np.float32)) for x in (q, k, v))).numpy()
    assert orc.max_abs_err(t, orc.attention_np(q, k, v)) < 2e-6


@pytest.mark.parametrize("n,m", [(1, 1), (5, 64), (64, 5)])
def test_oracle_known_answers(n, m):
    """Analytic KATs."""
    rng = np.random.default_rng(0)
    q = rng.standard_normal((1, 4, n, 64)).astype(np.float32)
    v = rng.standard_normal((1, 4, m, 64)).astype(np.float32)
    # all keys equal -> uniform weights -> O = mean(V)
    k = np.broadcast_to(rng.standard_normal((1, 4, 1, 64)).astype(np.float32), (1, 4, m, 64))
    o = orc.attention_np(q, k, v)
    np.testing.assert_allclose(o, np.broadcast_to(v.astype(np.float64).mean(axis=2, keepdims=True), o.shape), atol=1e-9)
    # V constant -> O = that constant
    k2 = rng.standard_normal((1, 4, m, 64)).astype(np.float32)
    vc = np.broadcast_to(v[:, :, :1], v.shape)
    np.testing.assert_allclose(orc.attention_np(q, k2, vc), np.broadcast_to(v[:, :, :1], o.shape), atol=1e-9)
    # permuting KV rows leaves O unchanged
    perm = rng.permutation(m)
    np.testing.assert_allclose(orc.attention_np(q, k2, v), orc.attention_np(q, k2[:, :, perm], v[:, :, perm]), atol=1e-9)
    # one dominant key -> O ~ its V row
    if m > 1:
        kd = np.zeros((1, 4, m, 64), np.float32)
        kd[:, :, 2 % m] = 100.0 * np.sign(q[:, :, 0])
        od = orc.attention_np(q[:, :, :1], kd, v)
        np.testing.assert_allclose(od[:, :, 0], v[:, :, 2 % m], atol=1e-9)


def test_work_formulas():
    assert orc.flops(1, 4, 1024, 1024) == pytest.approx(1.0737e9, rel=1e-3)
    assert orc.algorithmic_bytes(1, 4, 1024, 1024) == 2 * 1024 * 1024
    assert orc.algorithmic_bytes(1, 4, 1024, 1024, out_bytes=4) == 2.5 * 1024 * 1024
