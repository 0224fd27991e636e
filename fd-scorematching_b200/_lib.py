"""ctypes binding of libfdsm.so (include/fdsm.h).  No torch types cross this boundary: only raw
device pointers, sizes, scalars and the cudaStream_t handle.

There is NO CPU fallback: if the shared library is missing or a tensor is not a CUDA fp32 tensor
the call raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FDSM_LIB_PATH") or os.path.join(_HERE, "libfdsm.so")   # override: A/B runs of two builds

ROW_X, ROW_PLUS, ROW_MINUS = 1, 2, 4

_f32p = C.c_void_p
_i64 = C.c_int64
_i32 = C.c_int32
_u64 = C.c_uint64
_f32 = C.c_float
_stream = C.c_void_p


class StackArgs(C.Structure):
    """Mirror of `fdsm_stack_args` (include/fdsm.h)."""
    _fields_ = [
        ("x", _f32p), ("ldx", _i64),
        ("dir", _f32p), ("lddir", _i64),
        ("normalize_dir", _i32), ("dir_scale", _f32),
        ("noise", _f32p), ("ldnoise", _i64),
        ("sigma", _f32p), ("labels", C.c_void_p),
        ("perturb_input", _i32), ("dot_mode", _i32),
        ("out", _f32p), ("ldout", _i64),
        ("row_mask", _i32),
        ("v_out", _f32p), ("ldv", _i64),
        ("dot_out", _f32p),
        ("seed", _u64), ("offset", _u64), ("row0", _u64),
        ("B", _i64), ("D", _i64),
        ("out_lo", _f32p),
        ("offset_dev", C.c_void_p),
        ("pad_ones", _i32),
    ]


class JetArgs(C.Structure):
    """Mirror of `fdsm_jet_args` (include/fdsm.h)."""
    _fields_ = [
        ("x", _f32p), ("ldx", _i64), ("B", _i64),
        ("dir", _f32p), ("lddir", _i64),
        ("normalize_dir", _i32), ("dir_scale", _f32),
        ("seed", _u64), ("offset", _u64), ("row0", _u64),
        ("offset_dev", C.c_void_p),
        ("params", _f32p), ("workspace", _f32p),
        ("c1", _f32), ("scale", _f32), ("mean", _i32),
        ("loss", _f32p), ("grads", _f32p), ("accumulate", _i32),
        ("ev_chain_start", C.c_void_p), ("ev_chain_stop", C.c_void_p), ("trace", C.c_void_p),
        ("x_u8", C.c_void_p), ("x_u8_scale", _f32), ("x_noise_std", _f32), ("mode", _i32), ("update", C.c_void_p),
    ]


class JetUpdate(C.Structure):      # mirrors fdsm_jet_update
    _fields_ = [
        ("exp_avg", _f32p), ("exp_avg_sq", _f32p), ("lr", C.c_float), ("lr_dev", _f32p),
        ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("max_norm", C.c_float),
        ("step_counter", C.c_void_p), ("scratch", _f32p),
    ]


class GemmArgs(C.Structure):
    """Mirror of `fdsm_gemm_args` (include/fdsm.h)."""
    _fields_ = [
        ("a_hi", _f32p), ("a_lo", _f32p), ("lda", _i64), ("a_major", _i32),
        ("b_hi", _f32p), ("b_lo", _f32p), ("ldb", _i64), ("b_major", _i32),
        ("M", _i64), ("N", _i64), ("K", _i64),
        ("passes", _i32), ("splits", _i32), ("tile_n", _i32),
        ("out", _f32p), ("ldo", _i64), ("split_stride", _i64),
    ]


class MlpDesc(C.Structure):
    """Mirror of `fdsm_mlp_desc` (include/fdsm.h)."""
    _fields_ = [("in_dim", _i32), ("hidden", _i32), ("n_hidden", _i32), ("act", _i32), ("head", _i32), ("passes", _i32),
                ("presplit", _i32)]


ABI_VERSION = 11   # FDSM_ABI_VERSION of include/fdsm.h

# name -> (restype, argtypes); must list every symbol include/fdsm.h declares
PROTOTYPES = {
    "fdsm_version": (C.c_int, []),
    "fdsm_last_error_string": (C.c_char_p, []),
    "fdsm_device_supported": (C.c_int, []),
    "fdsm_stack": (C.c_int, [C.POINTER(StackArgs), _stream]),
    "fdsm_direction": (C.c_int, [_f32p, _i64, _f32p, _i64, _i64, _i64, _f32, _u64, _u64, _u64, _stream]),
    "fdsm_philox_normal": (C.c_int, [_f32p, _i64, _i64, _i64, _i32, _u64, _u64, _u64, _stream]),
    "fdsm_combine_energy_fwd": (C.c_int, [_f32p, _i64, _i32, _f32, _f32, _i32, _f32p, _f32p, _stream]),
    "fdsm_combine_energy_bwd": (C.c_int, [_f32p, _i64, _i32, _f32, _f32, _i32, _f32p, _f32p, _stream]),
    "fdsm_combine_dsm_fwd": (C.c_int, [_f32p, _f32p, _f32p, _i64, _f32, _f32, _f32, _f32p, _f32p, _stream]),
    "fdsm_combine_dsm_bwd": (C.c_int, [_f32p, _f32p, _f32p, _i64, _f32, _f32, _f32, _f32p, _f32p, _stream]),
    "fdsm_combine_score": (C.c_int, [_f32p, _i64, _f32p, _i64, _f32p, C.c_void_p, _f32, _i64, _i64, _f32, _i32,
                                     _f32p, _f32p, _f32p, _f32p, _i64, _stream]),
    "fdsm_scale_unless_one": (C.c_int, [_f32p, _i64, _f32p, _stream]),
    "fdsm_combine_score_static": (C.c_int, [_f32p, _i64, _f32p, _i64, _f32p, C.c_void_p, _f32, _i64, _i64, _f32, _i32,
                                            _f32p, _f32p, _f32p, _i64, _stream]),
    "fdsm_tf32_split": (C.c_int, [_f32p, _i64, _i64, _i64, _f32p, _f32p, _i64, _i64, _i32, _stream]),
    "fdsm_tc_gemm": (C.c_int, [C.POINTER(GemmArgs), _stream]),
    "fdsm_grad_sumsq": (C.c_int, [_f32p, _i64, _f32p, C.c_void_p, _stream]),
    "fdsm_adam_step": (C.c_int, [_f32p, _f32p, _f32p, _f32p, _i64, _f32, _f32p, _f32, _f32, _f32, _f32, _f32p, C.c_void_p,
                                 _stream]),
    "fdsm_publish_scalars": (C.c_int, [_f32p, _f32p, _i64, _stream]),
    "fdsm_stage_h2d": (C.c_int, [C.c_void_p, C.c_void_p, _i64, _stream, C.c_void_p, C.c_void_p, _stream]),
    "fdsm_publish_scalars_record": (C.c_int, [_f32p, _f32p, _i64, _stream, C.c_void_p]),
    "fdsm_mlp_sizes": (C.c_int, [C.POINTER(MlpDesc), _i64, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64)]),
    "fdsm_mlp_prepare": (C.c_int, [C.POINTER(MlpDesc), _i64, _f32p, _f32p, _stream]),
    "fdsm_mlp_forward": (C.c_int, [C.POINTER(MlpDesc), _f32p, _f32p, _i64, _f32p, _f32p, _f32p, _stream]),
    "fdsm_mlp_backward": (C.c_int, [C.POINTER(MlpDesc), _f32p, _f32p, _i64, _f32p, _f32p, _f32p, _f32p, _i32, _stream]),
    "fdsm_jet_supported": (C.c_int, [C.POINTER(MlpDesc)]),
    "fdsm_jet_sizes": (C.c_int, [C.POINTER(MlpDesc), _i64, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64)]),
    "fdsm_jet_fd_ssm": (C.c_int, [C.POINTER(MlpDesc), C.POINTER(JetArgs), _stream]),
    "fdsm_langevin_step": (C.c_int, [_f32p, _i64, _f32p, _i64, _f32p, _i64, _f32p, _i64, _i64, _i64, _f32, _f32, _u64, _u64, _u64,
                                     _stream]),
    "fdsm_ema_update": (C.c_int, [_f32p, _f32p, _i64, _f32, _f32, _stream]),
    "fdsm_inorm_plus_stats_floats": (C.c_int64, [_i64, _i64]),
    "fdsm_inorm_plus_launches": (C.c_int64, [_i64, _i64, _i64, _i32, _i32, _i32]),
    "fdsm_inorm_plus_fwd": (C.c_int, [_f32p, _i64, _i64, _i64, _f32p, _f32p, _f32p, _i64, _i32, _f32p, _f32p, _stream]),
    "fdsm_inorm_plus_bwd": (C.c_int, [_f32p, _f32p, _f32p, _i64, _i64, _i64, _f32p, _f32p, _i64, _i32, _f32p, _f32p, _f32p, _f32p,
                                      _f32p, _f32p, _stream]),
    "fdsm_peer_buffer_bytes": (C.c_int64, [_i64, _i32]),
    "fdsm_peer_allreduce": (C.c_int, [_f32p, _f32p, _i64, C.POINTER(C.c_void_p), _i32, _i32, _f32, _f32p, C.c_void_p,
                                      C.c_void_p, _stream]),
    "fdsm_mlp_backward_fd_ssm": (C.c_int, [C.POINTER(MlpDesc), _f32p, _f32p, _i64, _f32p, _f32p, _f32p, _f32, _f32, _i32,
                                           _f32p, _f32p, _i32, _stream]),
    "fdsm_mlp_backward_fd_dsm": (C.c_int, [C.POINTER(MlpDesc), _f32p, _f32p, _i64, _f32p, _f32p, _f32p, _f32p, _f32p, _f32,
                                           _f32, _f32, _f32p, _f32p, _i32, _stream]),
}

_lib = None


class FdsmError(RuntimeError):
    pass


def load():
    """Load libfdsm.so (once) and bind every prototype.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("FDSM_LIB", LIB_PATH)     # A/B builds of the same ABI (profiling); default: the in-tree library
    if not os.path.exists(path):
        raise FdsmError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C fd-scorematching_b200/csrc`).  There is no CPU/PyTorch fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype, fn.argtypes = res, args
    if lib.fdsm_version() != ABI_VERSION:
        raise FdsmError(f"libfdsm ABI version {lib.fdsm_version()} != {ABI_VERSION} (stale build: run __graft_entry__.build())")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().fdsm_last_error_string()
        raise FdsmError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
