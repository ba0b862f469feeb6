"""ctypes binding of libb2train.so (the C ABI declared in include/b2train.h).

There is no CPU or PyTorch fallback behind these calls: if the shared object is missing the import
fails, and every wrapper raises `B2Error` when the library reports an error.
"""
import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2_LIB: load another build of the same library (A/B measurements of a kernel change); default is the in-tree build
LIB_PATH = os.environ.get("B2_LIB") or os.path.join(_HERE, "libb2train.so")

F32, BF16 = 0, 1


class B2Error(RuntimeError):
    pass


class AdamWGroup(ctypes.Structure):
    """Mirror of B2AdamWGroup in include/b2train.h."""
    _fields_ = [("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("weight_decay", c_double), ("step", c_longlong)]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python gke-ray-train_b200/build.py` "
        "(there is no fallback implementation)")

lib = ctypes.CDLL(LIB_PATH)

_SIGS = {
    "b2_version": (c_int, []),
    "b2_last_error": (c_char_p, []),
    "b2_device_count": (c_int, []),
    "b2_adamw": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_double, c_double,
                         c_double, c_double, c_double, c_longlong, c_void_p, c_void_p]),
    "b2_adamw_grouped": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_void_p,
                                 c_void_p, c_int, c_void_p, c_void_p]),
    "b2_sumsq_scratch_bytes": (c_size_t, []),
    "b2_sumsq": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "b2_sumsq_static_max_elems": (c_size_t, []),
    "b2_sumsq_static": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_int, c_void_p]),
    "b2_clip_coef": (c_int, [c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "b2_scale": (c_int, [c_void_p, c_size_t, c_int, c_void_p, c_void_p]),
    "b2_ce_rows_fwd": (c_int, [c_void_p, c_int, c_longlong, c_longlong, c_longlong, c_void_p, c_int, c_longlong,
                               c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2_ce_rows_bwd": (c_int, [c_void_p, c_int, c_longlong, c_longlong, c_longlong, c_void_p, c_int, c_longlong,
                               c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2_lmhead_ce_workspace_bytes": (c_size_t, [c_longlong]),
    "b2_lmhead_ce_fwd": (c_int, [c_void_p, c_longlong, c_void_p, c_longlong, c_void_p, c_void_p, c_longlong, c_int,
                                 c_int, c_int, c_longlong, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p]),
    "b2_lmhead_ce_fwd_keep": (c_int, [c_void_p, c_longlong, c_void_p, c_longlong, c_void_p, c_void_p, c_longlong, c_int,
                                      c_int, c_int, c_longlong, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_longlong, c_void_p]),
    "b2_ce_dlogits_inplace": (c_int, [c_void_p, c_longlong, c_void_p, c_longlong, c_longlong, c_int, c_int, c_longlong,
                                      c_void_p, c_void_p, c_void_p]),
    "b2_lmhead_ce_bwd_dlogits": (c_int, [c_void_p, c_longlong, c_void_p, c_longlong, c_void_p, c_void_p, c_longlong,
                                         c_longlong, c_int, c_int, c_int, c_longlong, c_int, c_void_p, c_void_p,
                                         c_void_p, c_longlong, c_void_p]),
    "b2_lmhead_ce_bwd": (c_int, [c_void_p, c_longlong, c_void_p, c_longlong, c_void_p, c_void_p, c_longlong, c_longlong,
                                 c_int, c_int, c_int, c_longlong, c_int, c_void_p, c_void_p, c_void_p, c_longlong, c_int,
                                 c_void_p, c_longlong, c_void_p, c_int, c_longlong, c_int, c_void_p, c_int, c_void_p,
                                 c_size_t, c_void_p]),
    "b2_gemm_bf16": (c_int, [c_void_p, c_int, c_longlong, c_void_p, c_int, c_longlong, c_void_p, c_int, c_longlong,
                             c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "b2_gemm_workspace_bytes": (c_size_t, [c_int, c_int]),
    "b2_land_copy": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "b2_gather_windows": (c_int, [c_void_p, c_longlong, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2_peer_flags_bytes": (c_size_t, []),
    "b2_symm_alloc": (c_int, [c_size_t, c_void_p, c_void_p]),
    "b2_symm_open": (c_int, [c_void_p, c_void_p]),
    "b2_symm_close": (c_int, [c_void_p]),
    "b2_symm_free": (c_int, [c_void_p]),
    "b2_peer_allreduce": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, c_size_t, c_int, c_float, ctypes.c_uint,
                                  c_int, c_int, c_int, c_void_p, c_void_p]),
    "b2_peer_allreduce_mc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_size_t, c_size_t, c_int, c_float,
                                     ctypes.c_uint, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b2_peer_allgather": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, c_size_t, c_int, ctypes.c_uint, c_int,
                                  c_void_p]),
    "b2_peer_slice": (c_int, [c_size_t, c_int, c_int, c_int, c_void_p, c_void_p]),
}
# measurement aids: exported only by a library built with B2_BUILD_PROBES=1 (load it with B2_LIB=...)
_PROBE_SIGS = {
    "b2_debug_peer_copy": (c_int, [c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, c_void_p]),
}

for _name, (_res, _args) in list(_SIGS.items()) + [kv for kv in _PROBE_SIGS.items() if hasattr(lib, kv[0])]:
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args

EXPORTS = tuple(_SIGS)

# number of kernels this process has enqueued through the library (bench.py reports it)
launch_count = 0
_LAUNCHES = {"b2_adamw": 1, "b2_adamw_grouped": 1, "b2_sumsq": 1, "b2_sumsq_static": 1, "b2_clip_coef": 1, "b2_scale": 1, "b2_ce_rows_fwd": 2,
             "b2_ce_rows_bwd": 1, "b2_lmhead_ce_fwd": 3, "b2_lmhead_ce_fwd_keep": 3, "b2_ce_dlogits_inplace": 1, "b2_lmhead_ce_bwd_dlogits": 1, "b2_lmhead_ce_bwd": 3, "b2_gather_windows": 1, "b2_land_copy": 1, "b2_gemm_bf16": 1,  # (+1 reduce launch when split-K kicks in)
             "b2_peer_allreduce": 1, "b2_peer_allreduce_mc": 1, "b2_peer_allgather": 1}


def call(name, *args):
    """Calls an int-returning entry point and raises B2Error with the library's message on failure."""
    global launch_count
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise B2Error(f"{name} failed ({rc}): {lib.b2_last_error().decode(errors='replace')}")
    launch_count += _LAUNCHES.get(name, 0)
    return rc


def dtype_code(torch_dtype):
    import torch
    if torch_dtype == torch.float32:
        return F32
    if torch_dtype == torch.bfloat16:
        return BF16
    raise B2Error(f"unsupported dtype {torch_dtype}: the kernels take float32 or bfloat16")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B2Error("b200_train kernels need CUDA tensors; there is no CPU path")
