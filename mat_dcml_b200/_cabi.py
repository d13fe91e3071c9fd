"""ctypes binding of libmatas_b200.so (include/matas_b200.h).  There is no fallback: if the
library is missing or a call fails, a MatasError is raised."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libmatas_b200.so")

MODE_SEMI, MODE_DISCRETE = 0, 1


class MatasError(RuntimeError):
    pass


class ModelCfg(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in
                ("n_agent", "obs_dim", "state_dim", "act_dim", "n_block", "n_embd", "n_head", "mode")]


class EnvCfg(ctypes.Structure):
    _fields_ = [("n_envs", ctypes.c_long), ("n_workers", ctypes.c_int), ("seed", ctypes.c_uint64),
                ("env_offset", ctypes.c_long), ("double_reset", ctypes.c_int), ("shannon", ctypes.c_int),
                ("fixed", ctypes.c_int), ("standalone", ctypes.c_int)]


_P = ctypes.c_void_p
_SIGNATURES = {
    "matas_last_error": (ctypes.c_char_p, []),
    "matas_version": (ctypes.c_int, []),
    "matas_policy_create": (ctypes.c_int, [ctypes.POINTER(ModelCfg), ctypes.POINTER(_P)]),
    "matas_policy_destroy": (ctypes.c_int, [_P]),
    "matas_policy_param_count": (ctypes.c_long, [_P]),
    "matas_policy_set_weights": (ctypes.c_int, [_P, _P, _P]),
    "matas_policy_get_values": (ctypes.c_int, [_P, _P, ctypes.c_long, _P, _P]),
    "matas_policy_get_actions": (ctypes.c_int, [_P, _P, _P, ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_uint64,
                                                ctypes.c_uint32, ctypes.c_long, _P, _P, _P, _P, _P, _P]),
    "matas_policy_evaluate": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_long, _P, _P, _P, _P]),
    "matas_policy_evaluate_train": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_long, _P, _P, _P, _P]),
    "matas_policy_evaluate_backward": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_long, _P, _P, _P, _P, _P]),
    "matas_policy_encode": (ctypes.c_int, [_P, _P, ctypes.c_long, _P, _P, _P]),
    "matas_policy_act_autoregressive": (ctypes.c_int, [_P, _P, _P, ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_uint64,
                                                       ctypes.c_uint32, ctypes.c_long, _P, _P, _P, _P, _P]),
    "matas_policy_act_autoregressive_static": (ctypes.c_int, [_P, _P, _P, ctypes.c_long, ctypes.c_int, ctypes.c_int,
                                                              ctypes.c_uint64, ctypes.c_uint32, ctypes.c_long, _P, _P, _P, _P, _P]),
    "matas_policy_act_parallel": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_long, _P, _P, _P]),
    "matas_policy_encode_slice": (ctypes.c_int, [_P, _P, ctypes.c_long, ctypes.c_long, ctypes.c_long, _P, _P]),
    "matas_policy_decode": (ctypes.c_int, [_P, _P, ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_uint32,
                                           ctypes.c_long, _P, _P, _P, _P, _P]),
    "matas_policy_timing": (ctypes.c_int, [_P, ctypes.c_int]),
    "matas_policy_decode_time": (ctypes.c_int, [_P, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_long)]),
    "matas_gather_rows": (ctypes.c_int, [ctypes.c_int, _P, _P, _P, _P, _P, ctypes.c_long, _P]),
    "matas_philox4x32_10": (ctypes.c_int, [_P, _P, ctypes.c_long, _P]),
    "matas_ppo_loss": (ctypes.c_int, [ctypes.c_long, _P, _P, _P, _P, _P, _P, _P, _P, _P, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P, _P, _P, _P]),
    "matas_fp32_peak": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(ctypes.c_double)]),
    "dcml_env_create": (ctypes.c_int, [ctypes.POINTER(EnvCfg), _P, ctypes.POINTER(_P)]),
    "dcml_env_destroy": (ctypes.c_int, [_P]),
    "dcml_env_reset": (ctypes.c_int, [_P, _P, _P, _P, _P]),
    "dcml_env_reset_binary": (ctypes.c_int, [_P, _P, _P, _P, _P]),
    "dcml_env_step": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "dcml_env_step_slice": (ctypes.c_int, [_P, _P, ctypes.c_long, ctypes.c_long, ctypes.c_int, _P, _P, _P, _P, _P, _P, _P, _P]),
    "dcml_env_set_context": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "matas_adam_clip_step": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_long, ctypes.c_double, ctypes.c_double, ctypes.c_double,
                                            ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_long, _P, _P]),
    "matas_gae": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_double, ctypes.c_double, ctypes.c_int, ctypes.c_long, _P, _P, _P]),
}

_lib = None


def exported_symbols():
    """Names every build of the library must export (tests check them against the header)."""
    return sorted(_SIGNATURES)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MatasError("%s not found: build it with `python -m mat_dcml_b200.build` "
                             "(there is no CPU fallback)" % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            try:
                fn = getattr(handle, name)
            except AttributeError:
                raise MatasError("libmatas_b200.so does not export %s; rebuild it" % name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise MatasError("libmatas_b200: %s (code %d)" % (lib().matas_last_error().decode(), rc))


def ptr(t):
    """Device pointer of a contiguous torch tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "C-ABI arguments must be contiguous"
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
