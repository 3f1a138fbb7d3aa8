"""GpuDCRTPolyParams — host mirror of src/poly/dcrt/gpu.rs:267-594 (GpuDCRTPolyParams +
GpuContext) and of the CRT-basis arithmetic in src/poly/dcrt/params.rs:56-99.

The moduli are an INPUT to the GPU path (gpu.rs:567-594); `DCRTPolyParams.new`
obtains them from OpenFHE's GenCRTBasis, which is not available here, so
`standard_moduli` offers the descending prime chain that reproduces the one literal
basis in the reference tree ([131041, 131009] for n=4, 17 bits,
src/slot_transfer/bgg_poly_encoding_gpu.rs:1501).
"""
from __future__ import annotations

import ctypes as C

from . import ffi


def _is_prime(n: int) -> bool:
    if n < 2:
        return False
    small = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for p in small:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in small:
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def standard_moduli(ring_dimension: int, crt_depth: int, crt_bits: int) -> list[int]:
    """`crt_depth` largest primes below 2^crt_bits with p = 1 (mod 2n), descending."""
    step = 2 * ring_dimension
    c = ((1 << crt_bits) - 2) // step * step + 1
    out: list[int] = []
    while len(out) < crt_depth and c > step:
        if _is_prime(c):
            out.append(c)
        c -= step
    if len(out) != crt_depth:
        raise ValueError("not enough NTT-friendly primes below 2^crt_bits")
    return out


class GpuDCRTPolyParams:
    """ring dimension n, CRT moduli, gadget base 2^base_bits, one CUDA context.

    Mirrors GpuDCRTPolyParams::new_with_gpu (gpu.rs:567-594): `crt_bits` is the width of
    the widest modulus, `modulus_digits` = ceil(crt_bits / base_bits) * crt_depth
    (params.rs:56-58).
    """

    def __init__(self, ring_dimension: int, moduli, base_bits: int, gpu_ids=(0,), dnum: int = 0, library=None):
        n = int(ring_dimension)
        if n < 2 or n & (n - 1):
            raise ValueError("ring_dimension must be a power of 2")
        self._L = library or ffi.lib()
        self.ring_dimension = n
        self.log_ring = n.bit_length() - 1
        self.moduli = [int(q) for q in moduli]
        self.base_bits = int(base_bits)
        self.crt_depth = len(self.moduli)
        self.crt_bits = max(q.bit_length() for q in self.moduli)
        self.gpu_ids = tuple(int(g) for g in gpu_ids)
        mod = (C.c_uint64 * self.crt_depth)(*self.moduli)
        ids = (C.c_int * len(self.gpu_ids))(*self.gpu_ids)
        h = C.c_void_p()
        st = self._L.gpu_context_create(self.log_ring, self.crt_depth - 1, dnum, mod, self.crt_depth, ids, len(self.gpu_ids), C.byref(h))
        ffi.check_status(st, "gpu_context_create", self._L)
        self._ctx = h

    # -- PolyParams surface (src/poly/mod.rs:18-77) --------------------------------
    @property
    def modulus(self) -> int:
        q = 1
        for m in self.moduli:
            q *= m
        return q

    @property
    def digits_per_tower(self) -> int:
        return -(-self.crt_bits // self.base_bits)

    def modulus_digits(self) -> int:
        return self.digits_per_tower * self.crt_depth

    def ctx_raw(self):
        return self._ctx

    def get_n(self) -> int:
        n = C.c_int(0)
        ffi.check_status(self._L.gpu_context_get_N(self._ctx, C.byref(n)), "gpu_context_get_N", self._L)
        return n.value

    def stream(self) -> int:
        """cudaStream_t all work of this context is ordered on (extension)."""
        return int(self._L.gpu_context_stream(self._ctx) or 0)

    def synchronize(self) -> None:
        ffi.check_status(self._L.gpu_context_synchronize(self._ctx), "gpu_context_synchronize", self._L)

    def timer(self) -> "GpuTimer":
        return GpuTimer(self)

    def close(self) -> None:
        if getattr(self, "_ctx", None):
            self._L.gpu_context_destroy(self._ctx)
            self._ctx = None

    def __del__(self):  # Drop for GpuContext, gpu.rs:695-705
        try:
            self.close()
        except Exception:
            pass


class GpuTimer:
    """CUDA-event stopwatch on the context stream (extension gpu_timer_*)."""

    def __init__(self, params: GpuDCRTPolyParams):
        self._p = params
        self._L = params._L
        h = C.c_void_p()
        ffi.check_status(self._L.gpu_timer_create(params.ctx_raw(), C.byref(h)), "gpu_timer_create", self._L)
        self._h = h

    def start(self):
        ffi.check_status(self._L.gpu_timer_record(self._p.ctx_raw(), self._h, 0), "gpu_timer_record", self._L)

    def stop(self):
        ffi.check_status(self._L.gpu_timer_record(self._p.ctx_raw(), self._h, 1), "gpu_timer_record", self._L)

    def elapsed_ms(self) -> float:
        ms = C.c_float(0)
        ffi.check_status(self._L.gpu_timer_elapsed_ms(self._h, C.byref(ms)), "gpu_timer_elapsed_ms", self._L)
        return float(ms.value)

    def __del__(self):
        try:
            if self._h:
                self._L.gpu_timer_destroy(self._h)
                self._h = None
        except Exception:
            pass
