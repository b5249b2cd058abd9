"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by the product package).

CPU restatement of the reference's attention for the ``MHAHeadDim64`` hot path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import this module, and only as
the checker / the reported CPU baseline.

What it restates
----------------
``/root/reference/lightglue_pytorch_no_plugin/lightglue.py:82-85`` (``Attention.forward``)::

    query = query / 8
    qk = torch.matmul(query, key.permute(0, 1, 3, 2)).softmax(-1)
    qkv = torch.matmul(qk, value)

which is also what the plugin's eager twin computes
(``/root/reference/lightglue_pytorch_with_plugin/lightglue.py:46``,
``F.scaled_dot_product_attention`` with the default scale ``1/sqrt(64) = 1/8``).

The arithmetic under those four lines is PyTorch's CPU matmul/softmax (torch 2.11.0+cu128, not part
of /root/reference); the published algorithm is plain row-softmax attention, restated here in numpy
(float64 and float32) and, for timing, with the same torch calls.

Pinning: the reference ships NO tests or golden vectors for this path
(SURVEY.md §4, §8c).  The oracle is pinned instead against outputs of the reference itself:
``tests/golden/make_golden.py`` imports the unmodified reference modules in the build container and
commits their outputs (``Attention.forward`` and call-site captures of a seeded
``LightGlue(features=None)`` forward) as fixtures; ``tests/test_oracle.py`` checks this module
against them.
"""
from __future__ import annotations

import numpy as np

HEAD_DIM = 64
SCALE = 0.125  # 1 / sqrt(64); literal in the reference kernels (…fp32out.cu:468)


def attention_np(q, k, v, dtype=np.float64):
    """softmax(q k^T / 8) v, evaluated in ``dtype`` on the given values.

    q [B,H,N,64], k/v [B,H,M,64] (any float dtype; FP16 inputs are used at their exact values).
    Follows lightglue.py:82-85 step by step: scale q, scores, row softmax (max-subtracted, as
    torch.softmax does), weighted sum.
    """
    q = np.asarray(q, dtype=dtype)
    k = np.asarray(k, dtype=dtype)
    v = np.asarray(v, dtype=dtype)
    assert q.ndim == 4 and k.ndim == 4 and v.ndim == 4
    assert q.shape[-1] == HEAD_DIM and k.shape == v.shape and q.shape[:2] == k.shape[:2]
    qs = q / dtype(8)                                        # :82
    s = np.matmul(qs, np.swapaxes(k, -1, -2))                # :83 matmul
    s = s - s.max(axis=-1, keepdims=True)                    # :83 softmax(-1)
    p = np.exp(s)
    p = p / p.sum(axis=-1, keepdims=True)
    return np.matmul(p, v)                                   # :84


def attention_torch_cpu(q, k, v):
    """The same four lines with torch CPU ops in FP32 — the reference's CPU arithmetic itself.

    Used as the timed CPU baseline ("port"): identical calls to lightglue.py:82-85.
    """
    import torch

    q = q.float()
    k = k.float()
    v = v.float()
    q = q / 8
    qk = torch.matmul(q, k.permute(0, 1, 3, 2)).softmax(-1)
    return torch.matmul(qk, v)


def make_qkv(seed: int, batch: int, heads: int, n_q: int, n_kv: int, dtype=np.float16):
    """Seeded N(0,1) Q/K/V rounded to ``dtype`` (SURVEY.md §8d "D-attn")."""
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((batch, heads, n_q, HEAD_DIM), dtype=np.float32).astype(dtype)
    k = rng.standard_normal((batch, heads, n_kv, HEAD_DIM), dtype=np.float32).astype(dtype)
    v = rng.standard_normal((batch, heads, n_kv, HEAD_DIM), dtype=np.float32).astype(dtype)
    return q, k, v


def flops(batch: int, heads: int, n_q: int, n_kv: int) -> float:
    """Algorithmic FLOPs: 4*B*H*N*M*d (two GEMMs, 2 FLOP/MAC; softmax not counted)."""
    return 4.0 * batch * heads * n_q * n_kv * HEAD_DIM


def algorithmic_bytes(batch: int, heads: int, n_q: int, n_kv: int, in_bytes: int = 2, out_bytes: int = 2) -> float:
    """Algorithmic HBM bytes: read Q,K,V once, write O once."""
    return float(batch * heads * HEAD_DIM * (in_bytes * (n_q + 2 * n_kv) + out_bytes * n_q))


# tolerances stated by BASELINE.json's north_star
TOL_F16_MAX_ABS = 2e-3   # FP16 out: max-abs error vs the FP32/FP64 attention
TOL_F32_REL = 1e-3       # FP32 out: relative error (max-abs error / max-abs reference)


def max_abs_err(got, ref) -> float:
    return float(np.max(np.abs(np.asarray(got, dtype=np.float64) - np.asarray(ref, dtype=np.float64))))


def rel_err(got, ref) -> float:
    ref = np.asarray(ref, dtype=np.float64)
    return max_abs_err(got, ref) / float(np.max(np.abs(ref)))


def row_rel_err(got, ref) -> float:
    """The stricter reading of "relative" (reported beside rel_err, tests/test_attention_gpu.py): every output row's max-abs error
    over THAT row's max-abs reference value, maximised over rows — one normaliser per (batch, head, row) instead of one per tensor."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.max(np.abs(got - ref), axis=-1)
    den = np.max(np.abs(ref), axis=-1)
    return float(np.max(err / np.maximum(den, np.finfo(np.float64).tiny)))

