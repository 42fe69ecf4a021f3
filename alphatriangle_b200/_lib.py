"""ctypes binding of libat_b200.so (the C ABI in include/at_b200.h).

There is no CPU or PyTorch fallback behind this module: if the shared library is
missing, or an entry point fails, the caller gets an exception.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

AT_OK = 0
AT_E_INVALID, AT_E_CUDA, AT_E_STATE, AT_E_UNSUPPORTED = -1, -2, -3, -4

FLAG_NONFINITE_GRID = 1
FLAG_NONFINITE_OTHER = 2
FLAG_NONFINITE_POLICY_LOGITS = 4
FLAG_NONFINITE_VALUE_LOGITS = 8
FLAG_NONFINITE_POLICY_PROBS = 16
FLAG_PIPELINE_TIMEOUT = 0x80000000

PREC_FP32, PREC_BF16, PREC_F16 = 0, 1, 2
TENSOR_CORE_PRECISIONS = (PREC_BF16, PREC_F16)
PRECISION_NAMES = {PREC_FP32: "fp32", PREC_BF16: "bf16", PREC_F16: "f16"}
STATE_BYTES = 80
MAX_CONVS, MAX_MLP = 8, 4

ACTIVATIONS = {"ReLU": 1, "GELU": 2, "SiLU": 3, "Tanh": 4, "Sigmoid": 5}

LIB_NAME = "libat_b200.so"
LIB_PATH = Path(__file__).resolve().parent / LIB_NAME

# every symbol include/at_b200.h declares; tests check the built library exports all of them
EXPORTS = (
    "at_abi_version",
    "at_last_error",
    "at_extract_features",
    "at_model_describe",
    "at_engine_create",
    "at_engine_destroy",
    "at_engine_weight_floats",
    "at_engine_weight_entries",
    "at_engine_weight_entry",
    "at_engine_load_weights",
    "at_engine_publish_weights",
    "at_engine_weight_version",
    "at_engine_supports",
    "at_engine_forward",
    "at_engine_evaluate_packed",
    "at_engine_evaluate_packed_static",
    "at_engine_evaluate_host",
    "at_engine_submit_host",
    "at_engine_wait_host",
    "at_engine_last_launches",
    "at_engine_set_kernel_timing",
    "at_engine_kernel_ms",
    "at_engine_bf16_trunk",
    "at_engine_action_dim",
)


class LibraryError(RuntimeError):
    """The native library is missing or an entry point reported failure."""

    def __init__(self, message: str, code: int = 0):
        super().__init__(message)
        self.code = code


class EnvDesc(C.Structure):
    _fields_ = [("rows", C.c_int32), ("cols", C.c_int32), ("slots", C.c_int32)]


class ModelDesc(C.Structure):
    _fields_ = [
        ("env", EnvDesc),
        ("grid_channels", C.c_int32),
        ("n_conv", C.c_int32),
        ("conv_filters", C.c_int32 * MAX_CONVS),
        ("conv_kernel", (C.c_int32 * 2) * MAX_CONVS),
        ("conv_stride", (C.c_int32 * 2) * MAX_CONVS),
        ("conv_pad", (C.c_int32 * 2) * MAX_CONVS),
        ("n_res_blocks", C.c_int32),
        ("res_filters", C.c_int32),
        ("use_transformer", C.c_int32),
        ("t_dim", C.c_int32),
        ("t_heads", C.c_int32),
        ("t_layers", C.c_int32),
        ("t_ff", C.c_int32),
        ("n_shared", C.c_int32),
        ("shared_dims", C.c_int32 * MAX_MLP),
        ("n_policy", C.c_int32),
        ("policy_dims", C.c_int32 * MAX_MLP),
        ("n_value", C.c_int32),
        ("value_dims", C.c_int32 * MAX_MLP),
        ("num_atoms", C.c_int32),
        ("v_min", C.c_float),
        ("v_max", C.c_float),
        ("activation", C.c_int32),
        ("use_batch_norm", C.c_int32),
        ("other_dim", C.c_int32),
    ]


_lib = None


def _bind(lib: C.CDLL) -> None:
    vp, i32, i64, u32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32
    sig = {
        "at_abi_version": (i32, []),
        "at_last_error": (C.c_char_p, []),
        "at_extract_features": (i32, [C.POINTER(EnvDesc), i32, i32, vp, i64, vp, vp, vp]),
        "at_model_describe": (i64, [C.POINTER(ModelDesc), C.c_char_p, i64]),
        "at_engine_create": (i32, [C.POINTER(ModelDesc), i32, i64, C.POINTER(vp)]),
        "at_engine_destroy": (i32, [vp]),
        "at_engine_weight_floats": (i64, [vp]),
        "at_engine_weight_entries": (i32, [vp]),
        "at_engine_weight_entry": (i32, [vp, i32, C.POINTER(C.c_char_p), C.POINTER(i64), C.POINTER(i64)]),
        "at_engine_load_weights": (i32, [vp, vp, i64, i64, vp]),
        "at_engine_publish_weights": (i32, [vp, vp, i64, i64, vp]),
        "at_engine_weight_version": (i64, [vp]),
        "at_engine_supports": (i32, [vp, i32]),
        "at_engine_forward": (i32, [vp, i32, vp, vp, i64, vp, vp, vp, vp]),
        "at_engine_evaluate_packed": (i32, [vp, i32, vp, i64, vp, vp, vp, vp]),
        "at_engine_evaluate_packed_static": (i32, [vp, vp, i64, vp, vp, vp, vp]),
        "at_engine_evaluate_host": (i32, [vp, i32, vp, i64, vp, vp, vp, vp]),
        "at_engine_submit_host": (i32, [vp, i32, vp, i64, vp, vp, vp, i32, vp]),
        "at_engine_wait_host": (i32, [vp, i32]),
        "at_engine_last_launches": (i64, [vp]),
        "at_engine_bf16_trunk": (i32, [vp, vp, i64, vp, vp]),
        "at_engine_set_kernel_timing": (i32, [vp, i32]),
        "at_engine_kernel_ms": (i32, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "at_engine_action_dim": (i32, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def load() -> C.CDLL:
    """dlopen the in-tree library; raises LibraryError when it has not been built."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("AT_B200_LIB", LIB_PATH))
        if not path.exists():
            raise LibraryError(
                f"{path} not found: build it with `python __graft_entry__.py build` "
                "(nvcc, sm_100a). There is no fallback implementation."
            )
        lib = C.CDLL(str(path))
        _bind(lib)
        if lib.at_abi_version() != 1:
            raise LibraryError(f"{path}: ABI version {lib.at_abi_version()} != 1")
        _lib = lib
    return _lib


# include/at_b200_probe.h: the bring-up probes live in their own library (tests/test_gpu_probe.py, tools/mma_rate.py)
PROBE_LIB_PATH = Path(__file__).resolve().parent / "libat_b200_probe.so"
PROBE_EXPORTS = ("at_probe_last_error", "at_probe_umma", "at_probe_umma_ts", "at_probe_mma_rate", "at_probe_hog")
_probe = None


def load_probe() -> C.CDLL:
    global _probe
    if _probe is None:
        if not PROBE_LIB_PATH.exists():
            raise LibraryError(f"{PROBE_LIB_PATH} not found: build it with `python __graft_entry__.py build`")
        lib = C.CDLL(str(PROBE_LIB_PATH))
        vp, i32, i64, u32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32
        lib.at_probe_last_error.restype = C.c_char_p
        lib.at_probe_last_error.argtypes = []
        for name, args in (
            ("at_probe_umma_ts", [vp, vp, i64, i32, i32, i32, u32, u32, u32, u32, u32, u32, vp, vp, vp]),
            ("at_probe_mma_rate", [i32, i32, i32, i32, i32, vp, i32, vp, vp]),
            ("at_probe_hog", [i32, C.c_longlong, vp, vp]),
            ("at_probe_umma", [vp, i64, vp, i64, i32, i32, u32, u32, u32, u32, u32, u32, u32, u32, i32, vp, vp, vp]),
        ):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = i32, args
        _probe = lib
    return _probe


def check_probe(rc: int, what: str) -> None:
    if rc != AT_OK:
        msg = load_probe().at_probe_last_error()
        raise LibraryError(f"{what} failed ({rc}): {msg.decode('utf-8', 'replace') if msg else ''}", rc)


def last_error() -> str:
    msg = load().at_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, what: str) -> None:
    if rc != AT_OK:
        raise LibraryError(f"{what} failed ({rc}): {last_error()}", rc)


def _pair(v) -> tuple[int, int]:
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


def _resolve_padding(pad, kernel: tuple[int, int]) -> tuple[int, int]:
    if isinstance(pad, str):
        if pad == "valid":
            return 0, 0
        if pad == "same":
            if kernel[0] % 2 == 0 or kernel[1] % 2 == 0:
                raise ValueError("CONV_PADDING='same' needs odd kernel sizes")
            return kernel[0] // 2, kernel[1] // 2
        raise ValueError(f"unknown CONV_PADDING {pad!r}")
    return _pair(pad)


def env_desc(env_config) -> EnvDesc:
    return EnvDesc(int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS))


def model_desc(model_config, env_config) -> ModelDesc:
    mc = model_config
    d = ModelDesc()
    d.env = env_desc(env_config)
    d.grid_channels = mc.GRID_INPUT_CHANNELS
    if len(mc.CONV_FILTERS) > MAX_CONVS:
        raise ValueError(f"at most {MAX_CONVS} CONV_FILTERS entries are supported")
    d.n_conv = len(mc.CONV_FILTERS)
    for i, f in enumerate(mc.CONV_FILTERS):
        k = _pair(mc.CONV_KERNEL_SIZES[i])
        s = _pair(mc.CONV_STRIDES[i])
        p = _resolve_padding(mc.CONV_PADDING[i], k)
        d.conv_filters[i] = int(f)
        d.conv_kernel[i][0], d.conv_kernel[i][1] = k
        d.conv_stride[i][0], d.conv_stride[i][1] = s
        d.conv_pad[i][0], d.conv_pad[i][1] = p
    d.n_res_blocks = mc.NUM_RESIDUAL_BLOCKS
    d.res_filters = mc.RESIDUAL_BLOCK_FILTERS
    d.use_transformer = 1 if (mc.USE_TRANSFORMER and mc.TRANSFORMER_LAYERS > 0) else 0
    d.t_dim, d.t_heads, d.t_layers, d.t_ff = mc.TRANSFORMER_DIM, mc.TRANSFORMER_HEADS, mc.TRANSFORMER_LAYERS, mc.TRANSFORMER_FC_DIM
    for name, dims in (("shared", mc.FC_DIMS_SHARED), ("policy", mc.POLICY_HEAD_DIMS), ("value", mc.VALUE_HEAD_DIMS)):
        if len(dims) > MAX_MLP:
            raise ValueError(f"at most {MAX_MLP} hidden layers per MLP are supported")
        setattr(d, f"n_{name}", len(dims))
        arr = getattr(d, f"{name}_dims")
        for i, v in enumerate(dims):
            arr[i] = int(v)
    d.num_atoms = mc.NUM_VALUE_ATOMS
    d.v_min, d.v_max = float(mc.VALUE_MIN), float(mc.VALUE_MAX)
    d.activation = ACTIVATIONS[mc.ACTIVATION_FUNCTION]
    d.use_batch_norm = 1 if mc.USE_BATCH_NORM else 0
    d.other_dim = mc.OTHER_NN_INPUT_FEATURES_DIM
    return d


def weight_table(model_config, env_config) -> list[tuple[str, int, int]]:
    """(reference state_dict key, float offset, numel) for every float tensor, in image order. Needs no GPU."""
    lib = load()
    d = model_desc(model_config, env_config)
    need = lib.at_model_describe(C.byref(d), None, 0)
    if need < 0:
        raise LibraryError(f"at_model_describe failed ({need}): {last_error()}", int(need))
    buf = C.create_string_buffer(int(need) + 1)
    lib.at_model_describe(C.byref(d), buf, need + 1)
    rows = []
    for line in buf.value.decode().splitlines():
        name, off, numel = line.rsplit(" ", 2)
        rows.append((name, int(off), int(numel)))
    return rows
