"""Batched LightGlue pair matcher on a B200 — the CALLER of the attention hot path (SURVEY.md §8
rows e, f1, f2): what ``demo/lightglue_trt.cpp`` drives through the TensorRT engine, rebuilt so that
nothing but GEMMs, the attention kernel and four small glue kernels touch HBM.

Reference behaviour followed (``/root/reference/lightglue_pytorch_with_plugin/lightglue.py``):
``LightGlue.forward`` (:333-359) = 9 x [``SelfBlock`` (:137-151) on both images, ``CrossBlock``
(:188-206)] + ``MatchAssignment`` (:236-246) + ``filter_matches`` (:253-276).  ``input_dim ==
descriptor_dim == 256`` (SuperPoint), 4 heads x 64, no early stopping / pruning (:285-286).

B200-first layout (no translation of the reference's tensor choreography):

* both images of all ``B`` pairs live in ONE row-major FP16 activation matrix ``XC [R, 512]``,
  ``R = B*(N0+N1)``: columns 0..255 are the residual stream ``x``, columns 256..511 receive the
  attention output head-interleaved — which is exactly ``cat(x, context)``, the ffn input, so the
  ``torch.cat`` and every ``transpose/contiguous`` of the reference vanish;
* weights are re-packed once on the host (FP64 math, FP16 storage): ``Wqkv`` rows permuted from
  (head, dim, q|k|v) to (q|k|v, head, dim) so q, k, v are 256-column blocks of one GEMM output;
  ``to_qk``/``to_v`` stacked into one GEMM; ``out_proj``/``to_out`` folded into the first ffn
  matrix (``W1' = [W1x | W1m @ Wo]``); the final-projection scale ``256**-.25`` and the matchability
  row folded into one ``[264, 256]`` matrix; the constant second-ffn biases are carried symbolically
  (``x_true = x + c_l``) and folded into every consumer's bias, so the residual GEMM accumulates in place;
* GEMMs are plain cuBLAS calls (``torch.addmm`` / ``torch.bmm``); everything else is the hand-written
  sm_100a code behind ``include/lg_attn.h`` / ``include/lg_match.h``: strided attention (36 calls per
  pair, issued as 18 launches when both images carry the same keypoint count), rotary in place, LayerNorm+GELU in place, assignment + mutual-NN filter;
* the whole forward for a (B, N0, N1) shape is captured once into a CUDA graph and replayed.

There is no CPU path: tensors must be on a CUDA device and the C-ABI library must be built.
"""
from __future__ import annotations

import math
import os
from collections import OrderedDict
from dataclasses import dataclass

import torch

from . import _binding as B
from .ops import attention_strided

DIM = 256
HEADS = 4
HEAD_DIM = 64
FFN = 512
MD_COLS = 264  # 256 projected columns + 1 matchability logit, padded to a 16-byte multiple


# ------------------------------------------------------------------------------------------------
# weights
# ------------------------------------------------------------------------------------------------
def reference_init_state_dict(seed: int = 0, n_layers: int = 9) -> dict:
    """Random weights with the reference's parameter names, shapes and default initialisers, drawn in
    the order ``LightGlue.__init__`` creates them (lightglue.py:300-313) so that for a given
    ``torch.manual_seed`` the values equal those of ``LightGlue(features=None)`` (checked in
    tests/test_matcher_oracle.py when /root/reference is present).  There are no pretrained weights
    offline; this is the "random-init weights of that architecture" the bench runs on."""
    from torch import nn

    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    sd: dict[str, torch.Tensor] = {}

    def linear(name, fin, fout, bias=True):
        m = nn.Linear(fin, fout, bias=bias)
        sd[name + ".weight"] = m.weight.detach()
        if bias:
            sd[name + ".bias"] = m.bias.detach()
        return m

    wr = linear("posenc.Wr", 2, HEAD_DIM // 2, bias=False)
    nn.init.normal_(wr.weight.data, mean=0, std=1.0)  # gamma = 1 (lightglue.py:73-74)

    def ffn(prefix):
        linear(prefix + "ffn.0", 2 * DIM, 2 * DIM)
        sd[prefix + "ffn.1.weight"] = torch.ones(2 * DIM)
        sd[prefix + "ffn.1.bias"] = torch.zeros(2 * DIM)
        linear(prefix + "ffn.3", 2 * DIM, DIM)

    for i in range(n_layers):
        p = f"transformers.{i}.self_attn."
        linear(p + "Wqkv", DIM, 3 * DIM)
        linear(p + "out_proj", DIM, DIM)
        ffn(p)
        p = f"transformers.{i}.cross_attn."
        linear(p + "to_qk", DIM, DIM)
        linear(p + "to_v", DIM, DIM)
        linear(p + "to_out", DIM, DIM)
        ffn(p)
    for i in range(n_layers):
        linear(f"log_assignment.{i}.final_proj", DIM, DIM)
        linear(f"log_assignment.{i}.matchability", DIM, 1)
    torch.random.set_rng_state(gen_state)
    return sd


def load_checkpoint(path: str, n_layers: int = 9) -> dict:
    """A LightGlue ``.pth`` with the upstream key names, renamed the way the reference does
    (lightglue.py:322-328: ``self_attn.{i}`` -> ``transformers.{i}.self_attn`` …)."""
    sd = torch.load(path, map_location="cpu")
    for i in range(n_layers):
        sd = {k.replace(f"self_attn.{i}", f"transformers.{i}.self_attn"): v for k, v in sd.items()}
        sd = {k.replace(f"cross_attn.{i}", f"transformers.{i}.cross_attn"): v for k, v in sd.items()}
    return sd


@dataclass
class _Block:
    w_in_t: torch.Tensor    # [256, 768] (self: q|k|v) or [256, 512] (cross: qk|v), FP16, K-major rows
    b_in: torch.Tensor      # FP16
    w1_t: torch.Tensor      # [512, 512]: cat(x, context) -> hidden, out-projection folded in
    b1: torch.Tensor
    gamma: torch.Tensor     # FP32 [512]
    beta: torch.Tensor
    w2_t: torch.Tensor      # [512, 256]
    w1_nk: torch.Tensor = None   # [512 out, 512 in]: w1_t the other way round (nn.Linear layout), the fused ffn kernel's B operand
    w_in_nk: torch.Tensor = None # [768 | 512 out, 256 in]: w_in_t the other way round, the fused projection kernel's B operand


class MatcherWeights:
    """Host-side re-packing of a reference-format state dict (see the module docstring)."""

    def __init__(self, sd: dict, n_layers: int = 9, device="cuda:0"):
        self.n_layers = n_layers
        f64 = {k: v.detach().to(torch.float64).cpu() for k, v in sd.items()}
        dev = torch.device(device)

        def h(t):  # K-major FP16 matrix for x @ W^T
            return t.t().contiguous().to(dev, torch.float16)

        def vec(t, dtype=torch.float16):
            return t.contiguous().to(dev, dtype)

        self.wr = f64["posenc.Wr.weight"].to(dev, torch.float32).contiguous()
        # (head, dim, q|k|v) -> (q|k|v, head, dim): output row h*192 + d*3 + t moves to t*256 + h*64 + d
        perm = torch.arange(3 * DIM).reshape(HEADS, HEAD_DIM, 3).permute(2, 0, 1).reshape(-1)
        c = torch.zeros(DIM, dtype=torch.float64)  # accumulated second-ffn biases: x_true = x_stored + c
        self.blocks: list[tuple[_Block, _Block]] = []
        for i in range(n_layers):
            pair = []
            for kind in ("self_attn", "cross_attn"):
                p = f"transformers.{i}.{kind}."
                if kind == "self_attn":
                    w_in = f64[p + "Wqkv.weight"][perm]
                    b_in = f64[p + "Wqkv.bias"][perm]
                    wo, bo = f64[p + "out_proj.weight"], f64[p + "out_proj.bias"]
                else:
                    w_in = torch.cat([f64[p + "to_qk.weight"], f64[p + "to_v.weight"]], 0)
                    b_in = torch.cat([f64[p + "to_qk.bias"], f64[p + "to_v.bias"]], 0)
                    wo, bo = f64[p + "to_out.weight"], f64[p + "to_out.bias"]
                w1, b1 = f64[p + "ffn.0.weight"], f64[p + "ffn.0.bias"]
                w1x, w1m = w1[:, :DIM], w1[:, DIM:]
                w1f = torch.cat([w1x, w1m @ wo], 1)
                b1f = b1 + w1m @ bo + w1x @ c
                b_inf = b_in + w_in @ c
                pair.append(_Block(h(w_in), vec(b_inf), h(w1f), vec(b1f), vec(f64[p + "ffn.1.weight"], torch.float32),
                                   vec(f64[p + "ffn.1.bias"], torch.float32), h(f64[p + "ffn.3.weight"]),
                                   w1f.contiguous().to(dev, torch.float16), w_in.contiguous().to(dev, torch.float16)))
                c = c + f64[p + "ffn.3.bias"]
            self.blocks.append(tuple(pair))
        p = f"log_assignment.{n_layers - 1}."
        wf = torch.zeros(MD_COLS, DIM, dtype=torch.float64)
        bf = torch.zeros(MD_COLS, dtype=torch.float64)
        scale = DIM ** 0.25
        wf[:DIM] = f64[p + "final_proj.weight"] / scale
        bf[:DIM] = f64[p + "final_proj.bias"] / scale
        wf[DIM] = f64[p + "matchability.weight"][0]
        bf[DIM] = f64[p + "matchability.bias"][0]
        self.wf_t = h(wf)
        self.bf = vec(bf + wf @ c)
        self.c_final = vec(c)  # add to the stored residual stream to obtain lightglue_descriptors_{0,1}


def weights_blob(w: "MatcherWeights") -> bytes:
    """The packed weight blob of the native engine (include/lg_matcher.h; layout in csrc/lg_matcher.cu):
    header, posenc.Wr, per layer the self then the cross block (w_in_t, b_in, w1_t, b1, gamma, beta, w2_t),
    final projection, its bias, accumulated bias c — every array padded to 256 bytes."""
    import struct

    import numpy as np

    def pad(b: bytes) -> bytes:
        return b + b"\0" * (-len(b) % 256)

    def arr(t: torch.Tensor) -> bytes:
        return pad(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes())

    out = [pad(b"LGMATCH1" + struct.pack("<ii", w.n_layers, 0)), arr(w.wr)]
    for self_blk, cross_blk in w.blocks:
        for blk in (self_blk, cross_blk):
            out += [arr(blk.w_in_t), arr(blk.b_in), arr(blk.w1_t), arr(blk.b1), arr(blk.gamma), arr(blk.beta), arr(blk.w2_t)]
    out += [arr(w.wf_t), arr(w.bf), arr(w.c_final)]
    return b"".join(out)


class NativeMatcher:
    """ctypes front end of the native C++ engine (``lg_matcher_*``): same contract as ``PairMatcher.match`` but no
    torch in the call path — device pointers (or pinned host buffers) in, matches0 / mscores0 out."""

    def __init__(self, state_dict: dict | None = None, n_layers: int = 9, device: int = 0, threshold: float = 0.1,
                 seed: int = 0):
        import ctypes

        self._lib = B.load()
        sd = state_dict if state_dict is not None else reference_init_state_dict(seed, n_layers)
        blob = weights_blob(MatcherWeights(sd, n_layers, device="cpu"))
        if len(blob) != self._lib.lg_matcher_blob_bytes(n_layers):
            raise RuntimeError("weight blob size mismatch")
        self._h = ctypes.c_void_p()
        self.device = device
        B.check(self._lib.lg_matcher_create(ctypes.byref(self._h), device, n_layers, blob, len(blob), threshold),
                "lg_matcher_create")

    def plan(self, pairs: int, n0: int, n1: int) -> int:
        B.check(self._lib.lg_matcher_plan(self._h, pairs, n0, n1), "lg_matcher_plan")
        return self._lib.lg_matcher_launches(self._h)

    def match(self, kpts0: torch.Tensor, kpts1: torch.Tensor, desc0: torch.Tensor, desc1: torch.Tensor) -> dict:
        """CUDA tensors: kpts FP32 [B,N,2], desc FP16 [B,N,256] (contiguous).  Enqueued on the current stream."""
        b, n0, _ = kpts0.shape
        n1 = kpts1.shape[1]
        for t, dt in ((kpts0, torch.float32), (kpts1, torch.float32), (desc0, torch.float16), (desc1, torch.float16)):
            if not t.is_cuda or t.dtype != dt or not t.is_contiguous():
                raise ValueError("NativeMatcher.match: contiguous CUDA tensors, kpts float32 and desc float16")
        m0 = torch.empty(b, n0, dtype=torch.int32, device=kpts0.device)
        ms = torch.empty(b, n0, dtype=torch.float32, device=kpts0.device)
        with torch.cuda.device(kpts0.device):
            B.check(self._lib.lg_matcher_match(self._h, kpts0.data_ptr(), kpts1.data_ptr(), desc0.data_ptr(), desc1.data_ptr(),
                                               b, n0, n1, m0.data_ptr(), ms.data_ptr(),
                                               torch.cuda.current_stream(kpts0.device).cuda_stream), "lg_matcher_match")
        return {"matches0": m0, "mscores0": ms}

    def match_host(self, kpts0, kpts1, desc0, desc1):
        """numpy arrays / pinned CPU tensors in, numpy arrays out (H2D, forward, D2H, synchronise)."""
        import numpy as np

        k0, k1 = (np.ascontiguousarray(np.asarray(t), dtype=np.float32) for t in (kpts0, kpts1))
        d0, d1 = (np.ascontiguousarray(np.asarray(t), dtype=np.float16) for t in (desc0, desc1))
        b, n0, _ = k0.shape
        n1 = k1.shape[1]
        m0 = np.empty((b, n0), dtype=np.int32)
        ms = np.empty((b, n0), dtype=np.float32)
        B.check(self._lib.lg_matcher_match_host(self._h, k0.ctypes.data, k1.ctypes.data, d0.ctypes.data, d1.ctypes.data,
                                                b, n0, n1, m0.ctypes.data, ms.ctypes.data), "lg_matcher_match_host")
        return {"matches0": m0, "mscores0": ms}

    def submit_host(self, kpts0, kpts1, desc0, desc1, matches0, mscores0) -> None:
        """Asynchronous `match_host` on caller-owned (ideally pinned) C-contiguous arrays / CPU tensors: inputs are
        staged under the previous forward; `matches0` / `mscores0` are valid after `sync()`."""

        def ptr(x):
            return x.data_ptr() if isinstance(x, torch.Tensor) else x.ctypes.data

        b, n0 = int(kpts0.shape[0]), int(kpts0.shape[1])
        n1 = int(kpts1.shape[1])
        B.check(self._lib.lg_matcher_submit_host(self._h, ptr(kpts0), ptr(kpts1), ptr(desc0), ptr(desc1), b, n0, n1,
                                                 ptr(matches0), ptr(mscores0)), "lg_matcher_submit_host")

    def sync(self) -> None:
        B.check(self._lib.lg_matcher_sync(self._h), "lg_matcher_sync")

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.lg_matcher_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------
# glue-kernel wrappers (device pointers + current stream; include/lg_match.h)
# ------------------------------------------------------------------------------------------------
def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def posenc_table(kpts: torch.Tensor, wr: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """(cos, sin) of ``kpts @ wr^T``: kpts [rows, 2] FP32 -> [rows, 32, 2] FP32 (lightglue.py:77-86)."""
    if not kpts.is_cuda:
        raise RuntimeError("lg_match: tensors must live on a CUDA device (no CPU fallback)")
    kpts = kpts.contiguous()
    rows = kpts.numel() // 2
    if out is None:
        out = torch.empty(rows, 32, 2, dtype=torch.float32, device=kpts.device)
    B.check(B.load().lg_match_posenc(kpts.data_ptr(), wr.data_ptr(), out.data_ptr(), rows, _stream(kpts)),
            "lg_match_posenc")
    return out


def rotary_(packed: torch.Tensor, n_blocks: int, cs: torch.Tensor) -> torch.Tensor:
    """In-place rotary on the first ``n_blocks`` 256-column blocks of a packed FP16 [rows, ld] buffer."""
    if not packed.is_cuda or packed.dtype != torch.float16 or packed.dim() != 2 or packed.stride(1) != 1:
        raise ValueError("lg_match_rotary: expected a CUDA FP16 [rows, cols] buffer with contiguous rows")
    B.check(B.load().lg_match_rotary(packed.data_ptr(), packed.shape[0], packed.stride(0), n_blocks, cs.data_ptr(),
                                     _stream(packed)), "lg_match_rotary")
    return packed


def proj_rotary(x: torch.Tensor, w_nk: torch.Tensor, bias: torch.Tensor, n_rot_blocks: int, cs: torch.Tensor | None,
                out: torch.Tensor) -> torch.Tensor:
    """``out = x @ w_nk^T + bias`` with the first ``n_rot_blocks`` 256-column blocks rotated (``lg_match_proj_rotary``): x [rows, 256]
    FP16, w_nk [256 T, 256] FP16 contiguous (``nn.Linear.weight`` layout, T <= 3), bias FP16 [256 T], cs [rows, 32, 2] FP32 from
    ``posenc_table``, out [rows, 256 T] FP16.  Reference: lightglue_pytorch_with_plugin/lightglue.py:139-147, 190-193."""
    for t in (x, w_nk, bias, out) + ((cs,) if cs is not None else ()):
        if not t.is_cuda:
            raise RuntimeError("lg_match_proj_rotary: tensors must live on the CUDA device (no CPU path)")
    n_blocks = w_nk.shape[0] // DIM
    if (x.dtype != torch.float16 or out.dtype != torch.float16 or w_nk.dtype != torch.float16 or bias.dtype != torch.float16
            or x.dim() != 2 or x.shape[1] != DIM or x.stride(1) != 1 or w_nk.shape != (n_blocks * DIM, DIM) or not w_nk.is_contiguous()
            or bias.numel() != n_blocks * DIM or out.shape != (x.shape[0], n_blocks * DIM) or out.stride(1) != 1):
        raise ValueError("lg_match_proj_rotary: expected FP16 x [rows, 256], w [256 T, 256] contiguous, bias [256 T], out [rows, 256 T]")
    if n_rot_blocks and (cs is None or cs.dtype != torch.float32 or cs.numel() != x.shape[0] * 64 or not cs.is_contiguous()):
        raise ValueError("lg_match_proj_rotary: cs must be the contiguous FP32 [rows, 32, 2] table of posenc_table")
    B.check(B.load().lg_match_proj_rotary(x.data_ptr(), x.shape[0], x.stride(0), w_nk.data_ptr(), bias.data_ptr(), n_blocks,
                                          n_rot_blocks, cs.data_ptr() if cs is not None else None, out.data_ptr(), out.stride(0),
                                          _stream(x)), "lg_match_proj_rotary")
    return out


def ffn1_ln_gelu(x: torch.Tensor, w_nk: torch.Tensor, bias: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                 out: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """``out = GELU(LayerNorm(x @ w_nk^T + bias) * gamma + beta)`` as ONE launch (``lg_match_ffn1_ln_gelu``): x [rows, 512] FP16
    (row stride free), w_nk [512 out, 512 in] FP16 contiguous (``nn.Linear.weight`` layout), bias FP16 [512], gamma / beta FP32
    [512], out [rows, 512] FP16.  Reference: ffn[0..2], lightglue_pytorch_with_plugin/lightglue.py:131-136."""
    for t in (x, w_nk, bias, gamma, beta, out):
        if not t.is_cuda:
            raise RuntimeError("lg_match_ffn1_ln_gelu: tensors must live on the CUDA device (no CPU path)")
    if (x.dtype != torch.float16 or out.dtype != torch.float16 or w_nk.dtype != torch.float16 or bias.dtype != torch.float16
            or gamma.dtype != torch.float32 or beta.dtype != torch.float32):
        raise ValueError("lg_match_ffn1_ln_gelu: x / w / bias / out FP16, gamma / beta FP32")
    if (x.dim() != 2 or x.shape[1] != FFN or x.stride(1) != 1 or out.shape != x.shape or out.stride(1) != 1
            or w_nk.shape != (FFN, FFN) or not w_nk.is_contiguous() or bias.numel() != FFN or gamma.numel() != FFN or beta.numel() != FFN):
        raise ValueError("lg_match_ffn1_ln_gelu: expected x / out [rows, 512], w [512, 512] contiguous, 512-vectors")
    B.check(B.load().lg_match_ffn1_ln_gelu(x.data_ptr(), x.shape[0], x.stride(0), w_nk.data_ptr(), bias.data_ptr(), gamma.data_ptr(),
                                           beta.data_ptr(), eps, out.data_ptr(), out.stride(0), _stream(x)), "lg_match_ffn1_ln_gelu")
    return out


def layernorm_gelu_(h: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """In-place ``GELU(LayerNorm(h))`` on an FP16 [rows, 512] buffer (lightglue.py:131-136)."""
    if not h.is_cuda or h.dtype != torch.float16 or h.dim() != 2 or h.shape[1] != FFN or h.stride(1) != 1:
        raise ValueError("lg_match_layernorm_gelu: expected a CUDA FP16 [rows, 512] buffer")
    B.check(B.load().lg_match_layernorm_gelu(h.data_ptr(), h.shape[0], h.stride(0), gamma.data_ptr(), beta.data_ptr(),
                                             eps, _stream(h)), "lg_match_layernorm_gelu")
    return h


def assign(sim: torch.Tensor, z0: torch.Tensor, z1: torch.Tensor, threshold: float, want_scores: bool = False,
           out: dict | None = None):
    """Log double-softmax assignment + mutual-NN filter (lightglue.py:216-224, 253-276).

    sim [B,N0,N1] FP32 contiguous; z0 [B,N0] / z1 [B,N1]: FP16 views with a common element stride
    (column 256 of the packed final projection).  Returns matches0 (int32), mscores0, scores|None."""
    if not sim.is_cuda or sim.dtype != torch.float32 or not sim.is_contiguous():
        raise ValueError("lg_match_assign: sim must be a contiguous CUDA FP32 [B,N0,N1] tensor")
    b, n0, n1 = sim.shape
    zs = z0.stride(-1)
    if z1.stride(-1) != zs or z0.dtype != torch.float16 or z1.dtype != torch.float16:
        raise ValueError("lg_match_assign: z0 / z1 must be FP16 views with the same element stride")
    if (b > 1 and (z0.stride(0) != n0 * zs or z1.stride(0) != n1 * zs)):
        raise ValueError("lg_match_assign: z0 / z1 must be row-uniform views")
    lib = B.load()
    out = out or {}
    dev = sim.device
    ws_bytes = lib.lg_match_assign_workspace_bytes(b, n0, n1)
    ws = out.get("workspace")
    if ws is None or ws.numel() < ws_bytes:
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    m0 = out.get("matches0")
    if m0 is None:
        m0 = torch.empty(b, n0, dtype=torch.int32, device=dev)
    ms = out.get("mscores0")
    if ms is None:
        ms = torch.empty(b, n0, dtype=torch.float32, device=dev)
    scores = None
    if want_scores:
        scores = out.get("scores")
        if scores is None:
            scores = torch.empty_like(sim)
    B.check(lib.lg_match_assign(sim.data_ptr(), z0.data_ptr(), z1.data_ptr(), zs, b, n0, n1, float(threshold),
                                m0.data_ptr(), ms.data_ptr(), scores.data_ptr() if scores is not None else None,
                                ws.data_ptr(), ws_bytes, _stream(sim)), "lg_match_assign")
    return m0, ms, scores


# ------------------------------------------------------------------------------------------------
# the matcher
# ------------------------------------------------------------------------------------------------
class _Plan:
    """Device buffers + the captured graph for one (B, N0, N1) shape."""

    def __init__(self, b: int, n0: int, n1: int, dev, want_scores: bool):
        r = b * (n0 + n1)
        f16 = dict(dtype=torch.float16, device=dev)
        self.b, self.n0, self.n1, self.rows = b, n0, n1, r
        self.kpts = torch.empty(r, 2, dtype=torch.float32, device=dev)
        self.cs = torch.empty(r, 32, 2, dtype=torch.float32, device=dev)
        self.xc = torch.empty(r, FFN, **f16)
        self.proj = torch.empty(r, 3 * DIM, **f16)
        self.hid = torch.empty(r, FFN, **f16)
        self.md = torch.empty(r, MD_COLS, **f16)
        self.sim = torch.empty(b, n0, n1, dtype=torch.float32, device=dev)
        self.out = {
            "matches0": torch.empty(b, n0, dtype=torch.int32, device=dev),
            "mscores0": torch.empty(b, n0, dtype=torch.float32, device=dev),
            "workspace": torch.empty(B.load().lg_match_assign_workspace_bytes(b, n0, n1), dtype=torch.uint8, device=dev),
        }
        if want_scores:
            self.out["scores"] = torch.empty(b, n0, n1, dtype=torch.float32, device=dev)
        self.graph: torch.cuda.CUDAGraph | None = None
        self.launches = 0

    def stacked(self, t: torch.Tensor) -> torch.Tensor:
        """[2B, N, cols] view of both images' rows (only when N0 == N1): image 1 of pair b is entry B + b."""
        return t[: 2 * self.b * self.n0].unflatten(0, (2 * self.b, self.n0))

    def image_rows(self, t: torch.Tensor, img: int) -> torch.Tensor:
        """[B, N_img, cols] view of the rows of one image."""
        n = self.n0 if img == 0 else self.n1
        start = 0 if img == 0 else self.b * self.n0
        return t[start:start + self.b * n].unflatten(0, (self.b, n))


def _heads_view(t: torch.Tensor, col0: int) -> torch.Tensor:
    """[B, N, cols] -> [B, 4, N, 64] view of columns [col0, col0 + 256) (no copy)."""
    return t[..., col0:col0 + DIM].unflatten(-1, (HEADS, HEAD_DIM)).transpose(1, 2)


class PairMatcher:
    """``match(kpts0, kpts1, desc0, desc1)`` for ``B`` image pairs at once.

    kpts* [B,N,2] normalised keypoints (FP32), desc* [B,N,256] descriptors (FP16 or FP32), all on the
    matcher's device.  Returns ``{"matches0": int32 [B,N0], "mscores0": float32 [B,N0]}`` plus
    ``"scores"`` / ``"desc0"`` / ``"desc1"`` (the three engine outputs of
    export_onnx_with_plugin.py:197) when ``want_scores`` is set.  Results live in per-shape buffers
    that the next call with the same shape overwrites.
    """

    def __init__(self, state_dict: dict | None = None, n_layers: int = 9, device="cuda:0", threshold: float = 0.1,
                 seed: int = 0, use_graph: bool = True, max_plans: int = 8):
        if not torch.cuda.is_available():
            raise RuntimeError("PairMatcher needs a CUDA device (there is no CPU path)")
        B.load()
        self.device = torch.device(device)
        self.threshold = threshold
        self.use_graph = use_graph
        # LG_MATCH_FUSED_FFN=0: ffn[0..2] as library GEMM + lg_match_layernorm_gelu (round 2's path; kept for A/B runs)
        self.fused_ffn = os.environ.get("LG_MATCH_FUSED_FFN", "1") != "0"
        # LG_MATCH_FUSED_PROJ=0: input projections as library GEMM (+ lg_match_rotary) instead of lg_match_proj_rotary
        self.fused_proj = os.environ.get("LG_MATCH_FUSED_PROJ", "1") != "0"
        sd = state_dict if state_dict is not None else reference_init_state_dict(seed, n_layers)
        self.w = MatcherWeights(sd, n_layers, self.device)
        # shape plans, least recently used first.  The reference engine is dynamic (1..2048 keypoints per image, different for
        # every pair): an unbounded cache would grow by ~35 MB of buffers + one CUDA graph per distinct shape until OOM.
        self._plans: "OrderedDict[tuple, _Plan]" = OrderedDict()
        self.max_plans = int(max_plans)

    # -- one forward on the current stream (eager or under capture) ---------------------------------
    def _forward(self, p: _Plan, want_scores: bool) -> int:
        w = self.w
        x = p.xc[:, :DIM]
        launches = 0
        posenc_table(p.kpts, w.wr, p.cs)
        launches += 1
        for self_blk, cross_blk in w.blocks:
            # ---- self attention on both images (lightglue.py:137-151)
            if p.fused_proj_self:   # Wqkv with the rotary of q, k in the GEMM epilogue: one launch, no second pass over q | k
                proj_rotary(x, self_blk.w_in_nk, self_blk.b_in, 2, p.cs, p.proj)
                launches += 1
            else:
                torch.addmm(self_blk.b_in, x, self_blk.w_in_t, out=p.proj)
                rotary_(p.proj, 2, p.cs)
                launches += 2
            if p.n0 == p.n1:   # both images are one uniform [2B, N, .] batch: a single launch
                pj = p.stacked(p.proj)
                attention_strided(_heads_view(pj, 0), _heads_view(pj, DIM), _heads_view(pj, 2 * DIM),
                                  _heads_view(p.stacked(p.xc), DIM))
                launches += 1
            else:
                for img in (0, 1):
                    pj = p.image_rows(p.proj, img)
                    attention_strided(_heads_view(pj, 0), _heads_view(pj, DIM), _heads_view(pj, 2 * DIM),
                                      _heads_view(p.image_rows(p.xc, img), DIM))
                    launches += 1
            launches += self._ffn(p, self_blk)
            # ---- cross attention, both directions (lightglue.py:188-206)
            pc = p.proj.view(-1)[: p.rows * 2 * DIM].view(p.rows, 2 * DIM)
            if p.fused_proj_cross:
                proj_rotary(x, cross_blk.w_in_nk, cross_blk.b_in, 0, None, pc)
            else:
                torch.addmm(cross_blk.b_in, x, cross_blk.w_in_t, out=pc)
            launches += 1
            if p.n0 == p.n1:   # batch entry b attends to entry (b + B) mod 2B: both directions in one launch
                st = p.stacked(pc)
                attention_strided(_heads_view(st, 0), _heads_view(st, 0), _heads_view(st, DIM),
                                  _heads_view(p.stacked(p.xc), DIM), kv_batch_shift=p.b)
                launches += 1
            else:
                for img in (0, 1):
                    mine, other = p.image_rows(pc, img), p.image_rows(pc, 1 - img)
                    attention_strided(_heads_view(mine, 0), _heads_view(other, 0), _heads_view(other, DIM),
                                      _heads_view(p.image_rows(p.xc, img), DIM))
                    launches += 1
            launches += self._ffn(p, cross_blk)
        # ---- assignment (lightglue.py:236-246) + filter (:253-276)
        torch.addmm(w.bf, x, w.wf_t, out=p.md)
        md0, md1 = p.image_rows(p.md, 0), p.image_rows(p.md, 1)
        torch.bmm(md0[..., :DIM], md1[..., :DIM].transpose(1, 2), out_dtype=torch.float32, out=p.sim)
        assign(p.sim, md0[..., DIM], md1[..., DIM], self.threshold, want_scores, p.out)
        launches += 2 + B.load().lg_match_assign_launches(int(want_scores))
        return launches

    def _ffn(self, p: _Plan, blk: _Block) -> int:
        x = p.xc[:, :DIM]
        fused = self.fused_ffn and p.prefer_fused_ffn
        if fused:
            # cat(x, context) @ W1'^T + b1', LayerNorm, GELU: one tcgen05 kernel, the LayerNorm + GELU in its epilogue
            ffn1_ln_gelu(p.xc, blk.w1_nk, blk.b1, blk.gamma, blk.beta, p.hid)
        else:
            torch.addmm(blk.b1, p.xc, blk.w1_t, out=p.hid)      # library GEMM, then LayerNorm + GELU in place
            layernorm_gelu_(p.hid, blk.gamma, blk.beta)
        torch.addmm(x, p.hid, blk.w2_t, out=x)              # residual accumulate in place
        return 2 if fused else 3

    def plan(self, b: int, n0: int, n1: int, want_scores: bool = False) -> _Plan:
        key = (b, n0, n1, want_scores)
        p = self._plans.get(key)
        if p is not None:
            self._plans.move_to_end(key)
        if p is None:
            while len(self._plans) >= self.max_plans:
                # evicted buffers / graph may still be in use by a forward enqueued on some stream
                torch.cuda.synchronize(self.device)
                _, old = self._plans.popitem(last=False)
                old.graph = None
                del old
            with torch.cuda.device(self.device):
                p = _Plan(b, n0, n1, self.device, want_scores)
                B.check(B.load().lg_attn_hd64_init_device(self.device.index or 0), "lg_attn_hd64_init_device")
                B.check(B.load().lg_match_ffn_init_device(self.device.index or 0), "lg_match_ffn_init_device")
                p.prefer_fused_ffn = bool(B.load().lg_match_ffn1_prefer_fused(p.rows))   # few rows: library GEMM + LayerNorm/GELU kernel
                p.fused_proj_self = self.fused_proj and bool(B.load().lg_match_proj_prefer_fused(p.rows, 3, 2))
                p.fused_proj_cross = self.fused_proj and bool(B.load().lg_match_proj_prefer_fused(p.rows, 2, 0))
                if self.use_graph:
                    # warm up on a side stream (cuBLAS handles / workspaces), then capture once
                    p.kpts.zero_()
                    p.xc.zero_()
                    s = torch.cuda.Stream(self.device)
                    s.wait_stream(torch.cuda.current_stream(self.device))
                    with torch.cuda.stream(s):
                        self._forward(p, want_scores)
                    torch.cuda.current_stream(self.device).wait_stream(s)
                    torch.cuda.synchronize(self.device)
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        p.launches = self._forward(p, want_scores)
                    p.graph = g
            self._plans[key] = p
        return p

    def match(self, kpts0, kpts1, desc0, desc1, want_scores: bool = False) -> dict:
        for t in (kpts0, kpts1, desc0, desc1):
            if not t.is_cuda:
                raise RuntimeError("PairMatcher: inputs must live on the CUDA device (no CPU path)")
        b, n0, _ = kpts0.shape
        n1 = kpts1.shape[1]
        if desc0.shape != (b, n0, DIM) or desc1.shape != (b, n1, DIM) or kpts1.shape[0] != b:
            raise ValueError("PairMatcher: expected kpts [B,N,2] and desc [B,N,256] per image")
        p = self.plan(b, n0, n1, want_scores)
        with torch.cuda.device(self.device):
            p.image_rows(p.kpts, 0).copy_(kpts0)
            p.image_rows(p.kpts, 1).copy_(kpts1)
            p.image_rows(p.xc, 0)[..., :DIM].copy_(desc0)   # input_proj is the identity for 256-d descriptors
            p.image_rows(p.xc, 1)[..., :DIM].copy_(desc1)
            if p.graph is not None:
                p.graph.replay()
            else:
                p.launches = self._forward(p, want_scores)
        res = {"matches0": p.out["matches0"], "mscores0": p.out["mscores0"], "launches": p.launches + 4}
        if want_scores:
            res["scores"] = p.out["scores"]
            res["desc0"] = p.image_rows(p.xc, 0)[..., :DIM].float() + self.w.c_final.float()
            res["desc1"] = p.image_rows(p.xc, 1)[..., :DIM].float() + self.w.c_final.float()
        return res


def flops_per_pair(n0: int, n1: int, n_layers: int = 9) -> dict:
    """Algorithmic FLOPs of one pair (2 FLOP per MAC; reference op counts, no folding credit):
    attention 4*H*N*M*d per call, linear layers, assignment."""
    attn = n_layers * 4 * HEADS * HEAD_DIM * (n0 * n0 + n1 * n1 + 2 * n0 * n1)
    per_row = 2 * (DIM * 3 * DIM + DIM * DIM + FFN * FFN + FFN * DIM) + 2 * (2 * DIM * DIM + DIM * DIM + FFN * FFN + FFN * DIM)
    linear = n_layers * per_row * (n0 + n1) + 2 * (DIM * DIM + DIM) * (n0 + n1)
    sim = 2 * n0 * n1 * DIM
    return {"attention": attn, "linear": linear, "assignment": sim, "total": attn + linear + sim}
