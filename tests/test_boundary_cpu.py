"""CPU: the C-ABI library loads, exports every symbol include/lg_attn.h declares, and the TensorRT
plugin boundary (driven by the mock host) keeps the reference's contract.  No compute calls."""
import ctypes
import json
import re
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib(pkg):
    import __graft_entry__ as ge

    ge.build()
    return pkg.load()


def test_header_symbols_are_exported(pkg, lib):
    header = (ROOT / "include" / "lg_attn.h").read_text()
    declared = set(re.findall(r"\b(lg_attn_[a-z0-9_]+)\s*\(", header))
    assert declared == set(pkg.C_ABI_SYMBOLS)
    out = subprocess.run(["nm", "-D", "--defined-only", str(pkg._binding.lib_path())], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    for sym in declared:
        assert sym in exported, sym
        assert getattr(lib, sym) is not None
    assert lib.lg_attn_version() == 1
    # the shipped library carries no tuning switch (what-if arithmetic, phase traces): see include/lg_attn.h
    assert lib.lg_attn_build_flags() == 0


def test_glue_header_symbols_are_exported(pkg, lib):
    """include/lg_match.h (glue kernels either side of the path): every declared symbol is exported and bound."""
    header = (ROOT / "include" / "lg_match.h").read_text()
    declared = set(re.findall(r"^(?:int|size_t)\s+(lg_match_[a-z0-9_]+)\s*\(", header, flags=re.M))
    assert declared == set(pkg._binding.MATCH_ABI_SYMBOLS)
    out = subprocess.run(["nm", "-D", "--defined-only", str(pkg._binding.lib_path())], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    for sym in declared:
        assert sym in exported, sym
        assert getattr(lib, sym) is not None
    # argument errors are reported, never fatal, and need no GPU
    assert lib.lg_match_posenc(None, None, None, 4, None) == pkg._binding.EINVAL
    assert lib.lg_match_rotary(None, 4, 768, 2, None, None) == pkg._binding.EINVAL
    assert lib.lg_match_layernorm_gelu(None, 4, 512, None, None, 1e-5, None) == pkg._binding.EINVAL
    # the hand-written GEMM kernels of the matcher: bad pointers / strides / block counts are refused before any device call
    assert lib.lg_match_ffn1_ln_gelu(None, 4, 512, None, None, None, None, 1e-5, None, 512, None) == pkg._binding.EINVAL
    ok = ctypes.c_void_p(4096)
    assert lib.lg_match_ffn1_ln_gelu(ok, 4, 500, ok, ok, ok, ok, 1e-5, ok, 512, None) == pkg._binding.EINVAL        # ldx < 512
    assert lib.lg_match_ffn1_ln_gelu(ok, 4, 512, ok, ok, ok, ok, 1e-5, ctypes.c_void_p(4100), 512, None) == pkg._binding.EINVAL  # alignment
    assert lib.lg_match_proj_rotary(None, 4, 256, None, None, 3, 2, None, None, 768, None) == pkg._binding.EINVAL
    assert lib.lg_match_proj_rotary(ok, 4, 256, ok, ok, 4, 2, ok, ok, 1024, None) == pkg._binding.EINVAL            # more than 3 blocks
    assert lib.lg_match_proj_rotary(ok, 4, 256, ok, ok, 3, 2, None, ok, 768, None) == pkg._binding.EINVAL           # rotary without a table
    assert lib.lg_match_proj_rotary(ok, 4, 256, ok, ok, 3, 0, None, ok, 512, None) == pkg._binding.EINVAL           # ldo < 768
    # shape rules answer without a device (SM count defaults to a B200's 148): whole-row tiles need enough rows to fill the GPU,
    # except for the few-rows kernel, which needs an initialised device
    assert lib.lg_match_ffn1_prefer_fused(262144) == 1 and lib.lg_match_ffn1_prefer_fused(8192) == 0
    assert lib.lg_match_proj_prefer_fused(262144, 3, 2) == 1 and lib.lg_match_proj_prefer_fused(2048, 3, 2) == 0
    assert lib.lg_match_assign_workspace_bytes(0, 8, 8) == 0
    assert lib.lg_match_assign_workspace_bytes(2, 300, 200) >= 4 * 2 * (3 * 300 + 2 * 200)
    one = ctypes.c_void_p(256)
    assert lib.lg_match_assign(one, one, one, 1, 1, 8, 8, 0.1, one, one, None, None, 0, None) == pkg._binding.EWORKSPACE
    # the no-allocation variant refuses shapes beyond its fixed scratch before it touches any device
    assert lib.lg_match_assign_static(one, one, one, 1, 128, 1024, 1024, 0.1, one, one, None, None) == pkg._binding.ESTATIC_TOO_LARGE
    assert lib.lg_match_assign_launches(0) == 7 and lib.lg_match_assign_launches(1) == 8


def test_native_matcher_header_symbols_and_blob(pkg, lib):
    """include/lg_matcher.h (native C++ pair-matcher engine): symbols, blob size contract, blob validation."""
    header = (ROOT / "include" / "lg_matcher.h").read_text()
    declared = set(re.findall(r"^(?:int|size_t|void)\s+(lg_matcher_[a-z0-9_]+)\s*\(", header, flags=re.M))
    assert declared == set(pkg._binding.MATCHER_ABI_SYMBOLS)
    out = subprocess.run(["nm", "-D", "--defined-only", str(pkg._binding.lib_path())], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    assert declared <= exported
    m = pkg.matcher
    for n_layers in (1, 3):
        w = m.MatcherWeights(m.reference_init_state_dict(2, n_layers), n_layers, device="cpu")
        assert len(m.weights_blob(w)) == lib.lg_matcher_blob_bytes(n_layers)
    h = ctypes.c_void_p()
    blob = m.weights_blob(w)
    assert lib.lg_matcher_create(ctypes.byref(h), 0, 3, blob, len(blob) - 256, 0.1) == pkg._binding.EBLOB
    assert lib.lg_matcher_create(ctypes.byref(h), 0, 3, b"XXXXXXXX" + blob[8:], len(blob), 0.1) == pkg._binding.EBLOB
    assert lib.lg_matcher_create(ctypes.byref(h), 0, 0, blob, len(blob), 0.1) == pkg._binding.EINVAL
    assert lib.lg_matcher_launches(None) == 0


def test_library_is_sm100a_native(pkg):
    """The shipped kernel is tcgen05/TMA code for sm_100a (SASS mnemonics, B200_PROFILING.md)."""
    sass = subprocess.run(["cuobjdump", "-sass", str(pkg._binding.lib_path())], capture_output=True, text=True).stdout
    if not sass:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM", "STTM"):
        assert mnemonic in sass, mnemonic
    assert "HMMA." not in sass.replace("UTCHMMA", "")  # no legacy mma.sync path
    # per kernel: every hand-written GEMM-shaped kernel of the library (attention, fused ffn-up, projection) issues tcgen05.mma fed by
    # TMA, reads its accumulator with tcgen05.ld, and carries no local-memory spills in the one-CTA / pair ffn and attention kernels
    per_kernel, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            per_kernel[cur] = set()
        elif cur:
            m = re.match(r"\s+/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
            if m:
                per_kernel[cur].add(m.group(1))
    tc = {k: v for k, v in per_kernel.items() if any(n in k for n in ("lg_attn_hd64_fwd", "ffn1_ln_gelu", "proj_rotary_kernel"))}
    assert len(tc) >= 16, sorted(tc)   # 11 attention instantiations + 3 ffn + 2 projection
    for name, ops in tc.items():
        assert {"UTCHMMA", "UTMALDG", "LDTM", "UTCBAR"} <= ops, (name, sorted(ops))
        assert "HMMA" not in ops and "HGMMA" not in ops, name
    assert any("UTMASTG" in ops for name, ops in tc.items() if "ffn1_ln_gelu" in name)   # TMA stores out of the staging tiles


def test_cmake_build_drops_in(pkg, tmp_path):
    """The CMake flow of the reference's steps.txt (`cmake .. && make` in lightglue_attention_plugin/, CMakeLists.txt:11-12,76-89):
    csrc/CMakeLists.txt configures for sm_100a without CUTLASS / libtorch and yields the same library name exporting every C-ABI
    symbol and the plugin registration, with tcgen05 / TMA SASS inside."""
    import shutil

    if shutil.which("cmake") is None or shutil.which("ninja") is None:
        pytest.skip("cmake / ninja not installed")
    src = ROOT / "lightglue-with-flashattentionv2-tensorrt_b200" / "csrc"
    bld = tmp_path / "build"
    subprocess.run(["cmake", "-S", str(src), "-B", str(bld), "-G", "Ninja"], check=True, capture_output=True, text=True)
    subprocess.run(["cmake", "--build", str(bld), "-j", "8"], check=True, capture_output=True, text=True)
    lib_path = bld / "liblightglue_attention_plugin.so"
    assert lib_path.exists() and (bld / "lg_pair_bench").exists()
    out = subprocess.run(["nm", "-D", "--defined-only", str(lib_path)], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (\w+)", out))
    B = pkg._binding
    for sym in B.C_ABI_SYMBOLS + B.MATCH_ABI_SYMBOLS + B.MATCHER_ABI_SYMBOLS:
        assert sym in exported, sym
    elfs = subprocess.run(["cuobjdump", "-lelf", str(lib_path)], capture_output=True, text=True).stdout
    assert "sm_100a" in elfs and "sm_8" not in elfs and "sm_9" not in elfs
    # load it in a fresh process (this one already holds the in-tree build and its plugin registry): static registration
    # happened at load, the creator answers to the reference's plugin name / version
    child = (
        "import ctypes, json, sys\n"
        "lib = ctypes.CDLL(sys.argv[1])\n"
        "assert lib.lg_attn_version() == 1 and lib.lg_attn_build_flags() == 0\n"
        "lib.lg_mock_trt_describe.argtypes = [ctypes.c_char_p, ctypes.c_size_t]\n"
        "buf = ctypes.create_string_buffer(4096)\n"
        "assert lib.lg_mock_trt_describe(buf, 4096) > 0\n"
        "d = json.loads(buf.value.decode())\n"
        "assert (d['name'], d['version'], d['namespace'], d['n_creators']) == ('MHAHeadDim64', '1', '', 1), d\n"
        "print('ok')\n")
    r = subprocess.run([sys.executable, "-c", child, str(lib_path)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stderr[-800:]


def test_workspace_and_launch_counts(pkg, lib):
    assert lib.lg_attn_hd64_workspace_bytes(1, 4, 1024, 1024, pkg.F16_F16) == 0
    assert lib.lg_attn_hd64_workspace_bytes(1, 4, 1024, 1024, pkg.F16_F32) == 0
    assert lib.lg_attn_hd64_workspace_bytes(1, 4, 1024, 1000, pkg.F32_F32) == 0   # kFLOAT converts on the way into shared memory
    assert lib.lg_attn_hd64_workspace_bytes(1, 4, 0, 10, pkg.F32_F32) == 0
    assert lib.lg_attn_hd64_workspace_bytes(1, 4, 10, 10, 7) == 0
    assert lib.lg_attn_hd64_launches_per_call(pkg.F16_F16) == 1
    assert lib.lg_attn_hd64_launches_per_call(pkg.F16_F32) == 1
    assert lib.lg_attn_hd64_launches_per_call(pkg.F32_F32) == 1   # row a6: one launch (was convert pre-pass + attention)


def test_argument_errors_do_not_abort(pkg, lib):
    B = pkg._binding
    one = ctypes.c_void_p(16)
    assert lib.lg_attn_hd64_launch(None, one, one, one, 1, 4, 8, 8, 0, None, 0, None) == B.EINVAL
    assert lib.lg_attn_hd64_launch(one, one, one, one, 1, 4, 0, 8, 0, None, 0, None) == B.EINVAL
    assert lib.lg_attn_hd64_launch(one, one, one, one, 1, 4, 8, 8, 5, None, 0, None) == B.EINVAL
    assert lib.lg_attn_hd64_run_host(None, one, one, one, 1, 4, 8, 8, 0, 0) == B.EINVAL
    assert "workspace" in B.error_string(B.EWORKSPACE)
    assert B.error_string(0) == "success"


def test_plugin_contract_via_mock_host(lib):
    """Registry lookup + the build-time virtuals, vs the reference plugin's answers
    (lightglue_attention_plugin.cpp:19-20, :45-47, :77-94, :272-310, :363-367)."""
    lib.lg_mock_trt_describe.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    buf = ctypes.create_string_buffer(4096)
    n = lib.lg_mock_trt_describe(buf, 4096)
    assert n > 0
    d = json.loads(buf.value.decode())
    assert (d["name"], d["version"], d["namespace"]) == ("MHAHeadDim64", "1", "")
    assert (d["type"], d["plugin_version"]) == ("MHAHeadDim64", "1")
    assert d["n_creators"] == 1 and d["nb_fields"] == 0
    assert d["nb_outputs"] == 1
    assert d["serialization_size"] == 0 and d["serialize_untouched"]
    # slot 0: (float|half) & linear; slots 1-3: same type as slot 0 (= half here) & linear
    # order per slot: float/linear float/chw4 half/linear half/chw4 int8/linear int8/chw4
    assert d["format_table"] == "101000" + "001000" * 3
    assert d["format_follow_float"] == "10" * 3
    assert d["out_nb_dims"] == 4 and d["out_dims"] == [1, 4, 777, 64]  # O takes Q's dims
    assert d["out_type_half"] == 1 and d["out_type_float"] == 0        # DataType::kHALF / kFLOAT
    assert d["ws_half"] == 0
    assert d["ws_float"] == 0   # the reference asks TensorRT for a fixed 5 MiB (…plugin.cpp:96-112)
    assert d["clone_namespace"] == "ns_probe"


def test_python_operator_mirror(pkg):
    """Same op names as lightglue_pytorch_with_plugin/lightglue.py:16-46,102-115; no CPU fallback."""
    import torch

    assert pkg.ONNX_OP == "CustomOp::MHAHeadDim64"
    assert issubclass(pkg.MHAHeadDim64, torch.autograd.Function)
    assert hasattr(pkg.MHAHeadDim64, "symbolic")
    q = torch.zeros(1, 4, 8, 64)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.Attention()(q, q, q)


@pytest.mark.skipif(not Path("/root/reference/lightglue_pytorch_with_plugin/lightglue.py").exists(),
                    reason="the reference checkout only exists in the build container")
def test_bind_reference_model_routes_the_reference_call_sites_to_the_kernel(pkg):
    """bind_reference_model swaps the kernel in under the reference's own 36 call sites
    (lightglue_pytorch_with_plugin/lightglue.py:102-115 -> :148, :194, :198).  Checked here on CPU: after binding, the reference
    module's MHAHeadDim64.forward IS ours, its ONNX symbolic is untouched, and its Attention module no longer computes eagerly
    on CPU tensors — it raises the launcher's "no CPU fallback" error, i.e. the call really goes to the CUDA path."""
    import importlib
    import torch

    sys.path.insert(0, "/root/reference")
    try:
        ref_lg = importlib.import_module("lightglue_pytorch_with_plugin.lightglue")
    finally:
        sys.path.remove("/root/reference")
    ref_symbolic = ref_lg.MHAHeadDim64.symbolic
    eager_forward = ref_lg.MHAHeadDim64.forward
    q = torch.randn(1, 4, 8, 64)
    assert ref_lg.Attention()(q, q, q).shape == q.shape            # stock reference: eager SDPA on CPU
    try:
        pkg.bind_reference_model(ref_lg)
        assert ref_lg.MHAHeadDim64.forward is pkg.MHAHeadDim64.forward
        assert ref_lg.MHAHeadDim64.symbolic is ref_symbolic
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            ref_lg.Attention()(q, q, q)
        # the whole reference model now runs its attention through the binding as well
        model = ref_lg.LightGlue(features=None, n_layers=1).eval()
        data = {"keypoints0": torch.rand(1, 8, 2), "keypoints1": torch.rand(1, 6, 2),
                "descriptors0": torch.randn(1, 8, 256), "descriptors1": torch.randn(1, 6, 256)}
        with pytest.raises(RuntimeError, match="no CPU fallback"), torch.no_grad():
            model(data["keypoints0"], data["keypoints1"], data["descriptors0"], data["descriptors1"])
    finally:
        ref_lg.MHAHeadDim64.forward = eager_forward

