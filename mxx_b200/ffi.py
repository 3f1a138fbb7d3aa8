"""ctypes binding of libgpupoly.so — the Python twin of the reference's
`unsafe extern "C"` block (src/poly/dcrt/gpu.rs:69-240) and of `check_status`
(gpu.rs:259-263).  Prototypes are those of include/gpupoly.h.

There is no CPU fallback: if the library is missing or no CUDA device is visible the
first call raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgpupoly.so")

GPU_POLY_FORMAT_COEFF = 0  # gpu.rs:242
GPU_POLY_FORMAT_EVAL = 1  # gpu.rs:243

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u64p = C.POINTER(C.c_uint64)
intp = C.POINTER(C.c_int)
sizep = C.POINTER(C.c_size_t)
vpp = C.POINTER(C.c_void_p)
H = C.c_void_p  # opaque handle


class GpuRngSeed(C.Structure):  # gpu.rs:45-49
    _fields_ = [("words", C.c_uint64 * 4)]


# name -> (restype, argtypes); every symbol include/gpupoly.h declares
SIGNATURES = {
    "gpu_context_create": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, u64p, C.c_size_t, intp, C.c_size_t, vpp]),
    "gpu_context_destroy": (None, [H]),
    "gpu_context_get_N": (C.c_int, [H, intp]),
    "gpu_event_set_wait": (C.c_int, [H]),
    "gpu_event_set_destroy": (None, [H]),
    "gpu_device_synchronize": (C.c_int, []),
    "gpu_device_count": (C.c_int, [intp]),
    "gpu_device_mem_info": (C.c_int, [C.c_int, sizep, sizep]),
    "gpu_device_reset": (C.c_int, []),
    "gpu_last_error": (C.c_char_p, []),
    "gpu_pinned_alloc": (C.c_void_p, [C.c_size_t]),
    "gpu_pinned_free": (None, [C.c_void_p]),
    "gpu_matrix_create": (C.c_int, [H, C.c_int, C.c_size_t, C.c_size_t, C.c_int, vpp]),
    "gpu_matrix_destroy": (None, [H]),
    "gpu_matrix_copy": (C.c_int, [H, H]),
    "gpu_matrix_copy_block": (C.c_int, [H, H] + [C.c_size_t] * 6),
    "gpu_matrix_load_rns_batch": (C.c_int, [H, C.c_void_p, C.c_size_t, C.c_int, vpp]),
    "gpu_matrix_store_rns_batch": (C.c_int, [H, C.c_void_p, C.c_size_t, C.c_int, vpp]),
    "gpu_matrix_store_const_coeff_batch": (C.c_int, [H, u64p, C.c_size_t, vpp]),
    "gpu_matrix_store_compact_bytes": (C.c_int, [H, u8p, C.c_size_t, u16p, u16p, sizep]),
    "gpu_matrix_load_compact_bytes": (C.c_int, [H, u8p, C.c_size_t, C.c_uint16]),
    "gpu_matrix_add": (C.c_int, [H, H, H]),
    "gpu_matrix_add_block": (C.c_int, [H, H] + [C.c_size_t] * 6),
    "gpu_matrix_sub": (C.c_int, [H, H, H]),
    "gpu_matrix_mul": (C.c_int, [H, H, H]),
    "gpu_matrix_equal": (C.c_int, [H, H, intp]),
    "gpu_matrix_mul_scalar": (C.c_int, [H, H, H]),
    "gpu_matrix_ntt_all": (C.c_int, [H]),
    "gpu_matrix_intt_all": (C.c_int, [H]),
    "gpu_matrix_fill_gadget": (C.c_int, [H, C.c_uint32]),
    "gpu_matrix_fill_small_gadget": (C.c_int, [H, C.c_uint32]),
    "gpu_matrix_fill_small_decomposed_identity_chunk": (C.c_int, [H, H, C.c_size_t]),
    "gpu_matrix_decompose_base": (C.c_int, [H, C.c_uint32, H]),
    "gpu_matrix_decompose_base_small": (C.c_int, [H, C.c_uint32, H]),
    "gpu_matrix_gauss_samp_gq_arb_base": (C.c_int, [H, C.c_uint32, C.c_double, C.c_double, GpuRngSeed, H]),
    "gpu_matrix_create_p1_covariance_cache": (C.c_int, [H, H, H, C.c_double, C.c_double, C.c_double, vpp]),
    "gpu_matrix_destroy_p1_covariance_cache": (None, [H]),
    "gpu_matrix_sample_p1_full_cached": (C.c_int, [H, H, GpuRngSeed, H]),
    "gpu_matrix_sample_distribution": (C.c_int, [H, C.c_int, C.c_double, GpuRngSeed]),
    "gpu_matrix_sample_distribution_columns": (C.c_int, [H, C.c_int, C.c_double, GpuRngSeed, C.c_size_t, C.c_size_t]),
}
# entry points this library adds to the reference ABI (include/gpupoly.h "extensions")
EXTENSIONS = {
    "gpu_matrix_mul_decompose": (C.c_int, [H, H, H, C.c_uint32, C.c_int]),
    "gpu_matrix_mul_decompose_gathered": (C.c_int, [H, H, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, u64p, C.c_size_t, C.c_size_t,
                                                   C.c_size_t, C.c_size_t, C.c_uint32, C.c_int]),
    "gpu_matrix_intt_to": (C.c_int, [H, C.c_size_t, H]),
    "gpu_matrix_decompose_base_noalloc": (C.c_int, [H, C.c_uint32, H]),
    "gpu_matrix_add_scaled_block": (C.c_int, [H, H, H] + [C.c_size_t] * 6),
    "gpu_matrix_load_bigint_batch": (C.c_int, [H, u64p, C.c_size_t, C.c_int, vpp]),
    "gpu_matrix_store_bigint_batch": (C.c_int, [H, u64p, C.c_size_t, vpp]),
    "gpu_context_stream": (C.c_void_p, [H]),
    "gpu_thread_stream_override": (C.c_int, [C.c_void_p]),
    "gpu_matrix_device_info": (C.c_int, [H, vpp, intp, intp]),
    "gpu_kernel_launch_count": (C.c_uint64, []),
    "gpu_timer_create": (C.c_int, [H, vpp]),
    "gpu_timer_record": (C.c_int, [H, H, C.c_int]),
    "gpu_timer_elapsed_ms": (C.c_int, [H, C.POINTER(C.c_float)]),
    "gpu_timer_destroy": (None, [H]),
    "gpu_context_synchronize": (C.c_int, [H]),
}


def bind(path: str, extensions: bool = True) -> C.CDLL:
    """Load a library exporting the gpu_* ABI (ours, or the reference build used as a checker in tests)."""
    lib = C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    table = dict(SIGNATURES)
    if extensions:
        table.update(EXTENSIONS)
    for name, (res, args) in table.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if name == "gpu_device_reset":  # declared in Runtime.cuh:39, not bound by the Rust side
                continue
            raise
        fn.restype = res
        fn.argtypes = args
    return lib


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)"
            )
        _lib = bind(LIB_PATH)
    return _lib


class GpuError(RuntimeError):
    pass


def check_status(code: int, what: str, library: C.CDLL | None = None) -> None:
    """gpu.rs:259-263: non-zero status -> fail with the thread-local message."""
    if code != 0:
        msg = (library or lib()).gpu_last_error()
        raise GpuError(f"{what} failed: {msg.decode(errors='replace') if msg else 'unknown error'}")


def device_count() -> int:
    n = C.c_int(0)
    check_status(lib().gpu_device_count(C.byref(n)), "gpu_device_count")
    return n.value


def device_synchronize() -> None:
    check_status(lib().gpu_device_synchronize(), "gpu_device_synchronize")
