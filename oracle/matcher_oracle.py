"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by the product package).

CPU restatement (torch CPU, FP32 or FP64) of the reference's LightGlue forward — the CALLER of the
attention hot path — used to check the B200 pair matcher (e) and as the
timed CPU baseline of the pairs/s metric.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s CPU-baseline legs may import it.

What it restates (all under ``/root/reference/lightglue_pytorch_no_plugin/lightglue.py``; the
``_with_plugin`` twin is line-for-line the same apart from the attention call):

=========================  =======================================================================
``posenc``                 ``LearnableFourierPositionalEncoding.forward``
``rotary``                 ``SelfBlock.rotate_half`` / ``apply_cached_rotary_emb``
``self_block``             ``SelfBlock.forward`` (Wqkv head/dim/3 interleave, rotary on q and k)
``cross_block``            ``CrossBlock.forward`` (shared to_qk, both directions)
``log_assignment``         ``MatchAssignment.forward`` + ``sigmoid_log_double_softmax``
``filter_matches``         ``filter_matches`` (mutual nearest neighbour + threshold), returned in
                           the upstream LightGlue ``matches0`` / ``mscores0`` form
=========================  =======================================================================

It works directly on a reference-format ``state_dict`` (keys ``transformers.{i}.self_attn.Wqkv.weight``
…), so the very weights the reference model holds can be fed to it.

Pinning: the reference ships no tests or vectors.  ``tests/golden/make_golden_matcher.py``
runs the UNMODIFIED reference ``LightGlue(features=None)`` in the build container and commits its
outputs; ``tests/test_matcher_oracle.py`` checks this module against them (and, when
/root/reference is present, against the live reference model).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

HEADS = 4
HEAD_DIM = 64
DIM = 256


def posenc(kpts: torch.Tensor, wr: torch.Tensor):
    """kpts [B,N,2], wr [32,2] -> (cos, sin) each [B,1,N,64] with every angle repeated twice."""
    proj = kpts @ wr.t()                                  # Wr(x)
    cos, sin = torch.cos(proj), torch.sin(proj)
    cos = cos.repeat_interleave(2, dim=-1)[:, None]       # cat((emb, emb), -1).reshape(..., head_dim)
    sin = sin.repeat_interleave(2, dim=-1)[:, None]
    return cos, sin


def rotary(t: torch.Tensor, enc) -> torch.Tensor:
    """(t * cos) + (rotate_half(t) * sin), rotate_half: (a, b) -> (-b, a) on consecutive pairs."""
    cos, sin = enc
    pairs = t.unflatten(-1, (-1, 2))
    rot = torch.stack((-pairs[..., 1], pairs[..., 0]), dim=-1).flatten(-2)
    return t * cos + rot * sin


def attention(q, k, v):
    """lightglue.py:82-85."""
    p = torch.softmax((q / 8) @ k.transpose(-1, -2), dim=-1)
    return p @ v


def _ffn(sd, prefix, x, msg):
    h = F.linear(torch.cat((x, msg), -1), sd[prefix + "ffn.0.weight"], sd[prefix + "ffn.0.bias"])
    h = F.layer_norm(h, (2 * DIM,), sd[prefix + "ffn.1.weight"], sd[prefix + "ffn.1.bias"])
    h = F.gelu(h)
    return x + F.linear(h, sd[prefix + "ffn.3.weight"], sd[prefix + "ffn.3.bias"])


def _heads(t):  # [B,N,256] -> [B,4,N,64]
    return t.unflatten(-1, (HEADS, HEAD_DIM)).transpose(1, 2)


def self_block(sd, prefix, x, enc):
    b, n, _ = x.shape
    qkv = F.linear(x, sd[prefix + "Wqkv.weight"], sd[prefix + "Wqkv.bias"])
    qkv = qkv.reshape(b, n, HEADS, HEAD_DIM, 3).transpose(1, 2)
    q, k, v = qkv[..., 0], qkv[..., 1], qkv[..., 2]
    ctx = attention(rotary(q, enc), rotary(k, enc), v)
    ctx = ctx.transpose(1, 2).reshape(b, n, DIM)
    msg = F.linear(ctx, sd[prefix + "out_proj.weight"], sd[prefix + "out_proj.bias"])
    return _ffn(sd, prefix, x, msg)


def cross_block(sd, prefix, x0, x1):
    qk0 = _heads(F.linear(x0, sd[prefix + "to_qk.weight"], sd[prefix + "to_qk.bias"]))
    qk1 = _heads(F.linear(x1, sd[prefix + "to_qk.weight"], sd[prefix + "to_qk.bias"]))
    v0 = _heads(F.linear(x0, sd[prefix + "to_v.weight"], sd[prefix + "to_v.bias"]))
    v1 = _heads(F.linear(x1, sd[prefix + "to_v.weight"], sd[prefix + "to_v.bias"]))
    m0 = attention(qk0, qk1, v1).transpose(1, 2).flatten(-2)
    m1 = attention(qk1, qk0, v0).transpose(1, 2).flatten(-2)
    m0 = F.linear(m0, sd[prefix + "to_out.weight"], sd[prefix + "to_out.bias"])
    m1 = F.linear(m1, sd[prefix + "to_out.weight"], sd[prefix + "to_out.bias"])
    return _ffn(sd, prefix, x0, m0), _ffn(sd, prefix, x1, m1)


def log_assignment(sd, prefix, d0, d1):
    scale = DIM ** 0.25
    md0 = F.linear(d0, sd[prefix + "final_proj.weight"], sd[prefix + "final_proj.bias"]) / scale
    md1 = F.linear(d1, sd[prefix + "final_proj.weight"], sd[prefix + "final_proj.bias"]) / scale
    sim = md0 @ md1.transpose(1, 2)
    z0 = F.linear(d0, sd[prefix + "matchability.weight"], sd[prefix + "matchability.bias"])
    z1 = F.linear(d1, sd[prefix + "matchability.weight"], sd[prefix + "matchability.bias"])
    cert = F.logsigmoid(z0) + F.logsigmoid(z1).transpose(1, 2)
    return F.log_softmax(sim, 2) + F.log_softmax(sim, 1) + cert


def filter_matches(scores: torch.Tensor, th: float):
    """matches0 [B,N] (index into image 1 or -1) and mscores0 [B,N], from the log-assignment matrix."""
    max0, max1 = scores.max(2), scores.max(1)
    m0, m1 = max0.indices, max1.indices
    idx0 = torch.arange(m0.shape[1], device=m0.device)[None]
    mutual0 = idx0 == m1.gather(1, m0)
    mscores0 = torch.where(mutual0, max0.values.exp(), torch.zeros_like(max0.values))
    matches0 = torch.where(mscores0 > th, m0, torch.full_like(m0, -1))
    return matches0, mscores0


def forward(sd, kpts0, kpts1, desc0, desc1, n_layers: int = 9, dtype=torch.float32, threshold: float = 0.1):
    """Reference ``LightGlue.forward`` (input_dim == descriptor_dim) + ``filter_matches``.

    Returns dict(desc0, desc1, scores, matches0, mscores0).
    """
    sd = {k: v.to(dtype) for k, v in sd.items()}
    kpts0, kpts1, x0, x1 = (t.to(dtype) for t in (kpts0, kpts1, desc0, desc1))
    enc0, enc1 = posenc(kpts0, sd["posenc.Wr.weight"]), posenc(kpts1, sd["posenc.Wr.weight"])
    for i in range(n_layers):
        p = f"transformers.{i}."
        x0 = self_block(sd, p + "self_attn.", x0, enc0)
        x1 = self_block(sd, p + "self_attn.", x1, enc1)
        x0, x1 = cross_block(sd, p + "cross_attn.", x0, x1)
    scores = log_assignment(sd, f"log_assignment.{n_layers - 1}.", x0, x1)
    matches0, mscores0 = filter_matches(scores, threshold)
    return {"desc0": x0, "desc1": x1, "scores": scores, "matches0": matches0, "mscores0": mscores0}


def synth_pair(seed: int, n0: int, n1: int, width: int = 640, height: int = 480):
    """SuperPoint-shaped synthetic features (SURVEY.md §8d "D-feat"): distinct-ish integer pixels with a
    4-px border removed, normalised by ((x, y) - centre) / (max(w, h) / 2); L2-normalised 256-d descriptors.
    Returns kpts0 [1,n0,2], kpts1 [1,n1,2], desc0 [1,n0,256], desc1 [1,n1,256] (float32, CPU)."""
    g = torch.Generator().manual_seed(seed)

    def one(n):
        xs = torch.randint(4, width - 4, (n,), generator=g).float()
        ys = torch.randint(4, height - 4, (n,), generator=g).float()
        kp = (torch.stack([xs, ys], -1) - torch.tensor([width / 2.0, height / 2.0])) / (max(width, height) / 2.0)
        d = torch.randn(n, DIM, generator=g)
        return kp[None], (d / d.norm(dim=-1, keepdim=True))[None]

    k0, d0 = one(n0)
    k1, d1 = one(n1)
    return k0, k1, d0, d1
