"""ctypes binding of the C-ABI declared in ``include/lg_attn.h``.

The shared library is built in-tree by ``__graft_entry__.build()`` (or ``csrc/build.py``) as
``csrc/build/liblightglue_attention_plugin.so`` — the same file name the reference's CMake produces
(``/root/reference/lightglue_attention_plugin/CMakeLists.txt:76``) — so ``trtexec --plugins=`` /
``dlopen`` call sites keep working (``/root/reference/demo/lightglue_trt.cpp:15``).

There is no CPU fallback: if the library is missing, loading raises.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

_CSRC = Path(__file__).resolve().parent / "csrc"
LIB_NAME = "liblightglue_attention_plugin.so"
LIB_PATH = _CSRC / "build" / LIB_NAME

# io_dtype values of include/lg_attn.h
F16_F16 = 0
F32_F32 = 1
F16_F32 = 2

EINVAL = 100001
EWORKSPACE = 100002
EDRIVER = 100003
EARCH = 100004

# every symbol include/lg_attn.h declares (tests check the .so exports exactly these)
C_ABI_SYMBOLS = (
    "lg_attn_version",
    "lg_attn_build_flags",
    "lg_attn_debug_set_throughput_kernel",
    "lg_attn_debug_set_pdl",
    "lg_attn_debug_describe_launch",
    "lg_attn_hd64_init_device",
    "lg_attn_hd64_workspace_bytes",
    "lg_attn_hd64_launch",
    "lg_attn_hd64_launch_strided",
    "lg_attn_hd64_run_host",
    "lg_attn_hd64_launches_per_call",
    "lg_attn_error_string",
    "lg_attn_session_create",
    "lg_attn_session_submit",
    "lg_attn_session_submit_many",
    "lg_attn_session_sync",
    "lg_attn_session_destroy",
)
# every symbol include/lg_match.h declares
MATCH_ABI_SYMBOLS = (
    "lg_match_posenc",
    "lg_match_rotary",
    "lg_match_layernorm_gelu",
    "lg_match_ffn1_ln_gelu",
    "lg_match_ffn_init_device",
    "lg_match_ffn1_prefer_fused",
    "lg_match_debug_set_ffn_pair",
    "lg_match_proj_rotary",
    "lg_match_proj_prefer_fused",
    "lg_match_assign_workspace_bytes",
    "lg_match_assign",
    "lg_match_assign_static",
    "lg_match_assign_launches",
)

# every symbol include/lg_matcher.h declares
MATCHER_ABI_SYMBOLS = (
    "lg_matcher_create",
    "lg_matcher_blob_bytes",
    "lg_matcher_plan",
    "lg_matcher_match",
    "lg_matcher_match_host",
    "lg_matcher_submit_host",
    "lg_matcher_sync",
    "lg_matcher_launches",
    "lg_matcher_plans_resident",
    "lg_matcher_destroy",
)
ECUBLAS = 100010
EBLOB = 100011
ESTATIC_TOO_LARGE = 100012
ASSIGN_STATIC_MAX_BYTES = 4194304

_lib = None


class Strides(ctypes.Structure):
    """``lg_attn_strides``: element strides of the batch / head / row dimensions."""
    _fields_ = [("batch", ctypes.c_longlong), ("head", ctypes.c_longlong), ("row", ctypes.c_longlong)]


class LgAttnError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(f"{what} failed: code {code} ({error_string(code)})")
        self.code = code


def describe_launch(batch: int, heads: int, n_q: int, n_kv: int) -> str:
    """Which kernel / grid ``lg_attn_hd64_launch`` uses for this shape on the current device."""
    buf = ctypes.create_string_buffer(256)
    check(load().lg_attn_debug_describe_launch(batch, heads, n_q, n_kv, buf, 256), "lg_attn_debug_describe_launch")
    return buf.value.decode()


def lib_path() -> Path:
    override = os.environ.get("LG_ATTN_LIB")
    return Path(override) if override else LIB_PATH


def load() -> ctypes.CDLL:
    """Load the C-ABI library (once). Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        raise FileNotFoundError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the attention path)"
        )
    lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
    lib.lg_attn_version.restype = ctypes.c_int
    lib.lg_attn_version.argtypes = []
    lib.lg_attn_build_flags.restype = ctypes.c_int
    lib.lg_attn_build_flags.argtypes = []
    lib.lg_attn_debug_set_throughput_kernel.restype = ctypes.c_int
    lib.lg_attn_debug_set_throughput_kernel.argtypes = [ctypes.c_int]
    lib.lg_attn_debug_set_pdl.restype = ctypes.c_int
    lib.lg_attn_debug_set_pdl.argtypes = [ctypes.c_int]
    lib.lg_attn_debug_describe_launch.restype = ctypes.c_int
    lib.lg_attn_debug_describe_launch.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char_p, ctypes.c_size_t]
    lib.lg_attn_hd64_init_device.restype = ctypes.c_int
    lib.lg_attn_hd64_init_device.argtypes = [ctypes.c_int]
    lib.lg_attn_hd64_workspace_bytes.restype = ctypes.c_size_t
    lib.lg_attn_hd64_workspace_bytes.argtypes = [ctypes.c_int] * 5
    lib.lg_attn_hd64_launch.restype = ctypes.c_int
    lib.lg_attn_hd64_launch.argtypes = [
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
        ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
    ]
    lib.lg_attn_hd64_launch_strided.restype = ctypes.c_int
    lib.lg_attn_hd64_launch_strided.argtypes = [
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
        ctypes.POINTER(Strides), ctypes.POINTER(Strides), ctypes.POINTER(Strides), ctypes.POINTER(Strides),
        ctypes.c_int, ctypes.c_void_p,
    ]
    lib.lg_attn_hd64_run_host.restype = ctypes.c_int
    lib.lg_attn_hd64_run_host.argtypes = [
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
    ]
    lib.lg_attn_hd64_launches_per_call.restype = ctypes.c_int
    lib.lg_attn_hd64_launches_per_call.argtypes = [ctypes.c_int]
    lib.lg_attn_error_string.restype = ctypes.c_char_p
    lib.lg_attn_error_string.argtypes = [ctypes.c_int]
    lib.lg_attn_session_create.restype = ctypes.c_int
    lib.lg_attn_session_create.argtypes = [ctypes.POINTER(ctypes.c_void_p)] + [ctypes.c_int] * 7
    lib.lg_attn_session_submit.restype = ctypes.c_int
    lib.lg_attn_session_submit.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] * 3
    lib.lg_attn_session_submit_many.restype = ctypes.c_int
    lib.lg_attn_session_submit_many.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] + [ctypes.c_size_t] * 3 + [ctypes.c_int] * 4
    lib.lg_attn_session_sync.restype = ctypes.c_int
    lib.lg_attn_session_sync.argtypes = [ctypes.c_void_p]
    lib.lg_attn_session_destroy.restype = None
    lib.lg_attn_session_destroy.argtypes = [ctypes.c_void_p]
    P, LL, I, F = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_float
    lib.lg_match_posenc.restype = I
    lib.lg_match_posenc.argtypes = [P, P, P, LL, P]
    lib.lg_match_rotary.restype = I
    lib.lg_match_rotary.argtypes = [P, LL, I, I, P, P]
    lib.lg_match_layernorm_gelu.restype = I
    lib.lg_match_layernorm_gelu.argtypes = [P, LL, I, P, P, F, P]
    lib.lg_match_ffn1_ln_gelu.restype = I
    lib.lg_match_ffn1_ln_gelu.argtypes = [P, LL, I, P, P, P, P, F, P, I, P]
    lib.lg_match_ffn_init_device.restype = I
    lib.lg_match_ffn_init_device.argtypes = [I]
    lib.lg_match_ffn1_prefer_fused.restype = I
    lib.lg_match_ffn1_prefer_fused.argtypes = [LL]
    lib.lg_match_debug_set_ffn_pair.restype = I
    lib.lg_match_debug_set_ffn_pair.argtypes = [I]
    lib.lg_match_proj_rotary.restype = I
    lib.lg_match_proj_rotary.argtypes = [P, LL, I, P, P, I, I, P, P, I, P]
    lib.lg_match_proj_prefer_fused.restype = I
    lib.lg_match_proj_prefer_fused.argtypes = [LL, I, I]
    lib.lg_match_assign_workspace_bytes.restype = ctypes.c_size_t
    lib.lg_match_assign_workspace_bytes.argtypes = [I, I, I]
    lib.lg_match_assign.restype = I
    lib.lg_match_assign.argtypes = [P, P, P, LL, I, I, I, F, P, P, P, P, ctypes.c_size_t, P]
    lib.lg_match_assign_static.restype = I
    lib.lg_match_assign_static.argtypes = [P, P, P, LL, I, I, I, F, P, P, P, P]
    lib.lg_match_assign_launches.restype = I
    lib.lg_match_assign_launches.argtypes = [I]
    lib.lg_matcher_create.restype = I
    lib.lg_matcher_create.argtypes = [ctypes.POINTER(ctypes.c_void_p), I, I, P, ctypes.c_size_t, F]
    lib.lg_matcher_blob_bytes.restype = ctypes.c_size_t
    lib.lg_matcher_blob_bytes.argtypes = [I]
    lib.lg_matcher_plan.restype = I
    lib.lg_matcher_plan.argtypes = [P, I, I, I]
    lib.lg_matcher_match.restype = I
    lib.lg_matcher_match.argtypes = [P, P, P, P, P, I, I, I, P, P, P]
    lib.lg_matcher_match_host.restype = I
    lib.lg_matcher_match_host.argtypes = [P, P, P, P, P, I, I, I, P, P]
    lib.lg_matcher_submit_host.restype = I
    lib.lg_matcher_submit_host.argtypes = [P, P, P, P, P, I, I, I, P, P]
    lib.lg_matcher_sync.restype = I
    lib.lg_matcher_sync.argtypes = [P]
    lib.lg_matcher_launches.restype = I
    lib.lg_matcher_launches.argtypes = [P]
    lib.lg_matcher_plans_resident.restype = I
    lib.lg_matcher_plans_resident.argtypes = [P]
    lib.lg_matcher_destroy.restype = None
    lib.lg_matcher_destroy.argtypes = [P]
    _lib = lib
    return lib


def error_string(code: int) -> str:
    return load().lg_attn_error_string(int(code)).decode()


def check(code: int, what: str) -> None:
    if code != 0:
        raise LgAttnError(code, what)
