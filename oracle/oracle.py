"""ctypes/numpy front end of the CPU oracle (oracle/dcrt_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never by mxx_b200/.

Matrices are numpy uint64 arrays of shape (rows, cols, L, n) — the host layout
[poly row-major][limb][coeff] of src/matrix/gpu_dcrt_poly.rs:769-817.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("dcrt_oracle.c", "sampler_oracle.c", "dcrt_oracle.h")]
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        sig = {
            "orc_mulmod": (C.c_uint64, [C.c_uint64] * 3),
            "orc_powmod": (C.c_uint64, [C.c_uint64] * 3),
            "orc_invmod": (C.c_uint64, [C.c_uint64] * 2),
            "orc_barrett_mulmod_check": (C.c_uint64, [C.c_uint64] * 3),
            "orc_is_prime": (C.c_int, [C.c_uint64]),
            "orc_least_primitive_root": (C.c_uint64, [C.c_uint64]),
            "orc_psi": (C.c_uint64, [C.c_uint64, C.c_uint64]),
            "orc_prime_chain": (C.c_int, [C.c_uint64, C.c_uint32, C.c_uint32, u64p]),
            "orc_rand_residue": (C.c_uint64, [C.c_uint64] * 3),
            "orc_fill_uniform": (None, [u64p, C.c_size_t, C.c_uint32, C.c_size_t, u64p, C.c_uint64]),
            "orc_ntt_naive": (None, [u64p, u64p, C.c_size_t, C.c_uint64, C.c_uint64]),
            "orc_ntt_ref": (None, [u64p, C.c_size_t, C.c_uint64, C.c_uint64]),
            "orc_intt_ref": (None, [u64p, C.c_size_t, C.c_uint64, C.c_uint64]),
            "orc_negacyclic_schoolbook": (None, [u64p, u64p, u64p, C.c_size_t, C.c_uint64]),
            "orc_plan_create": (C.c_void_p, [C.c_size_t, u64p, C.c_uint32]),
            "orc_plan_destroy": (None, [C.c_void_p]),
            "orc_get_threads": (C.c_int, []),
            "orc_set_threads": (None, [C.c_int]),
            "orc_matrix_ntt": (None, [C.c_void_p, u64p, C.c_size_t]),
            "orc_matrix_intt": (None, [C.c_void_p, u64p, C.c_size_t]),
            "orc_matrix_mul": (None, [C.c_void_p, u64p, u64p, u64p] + [C.c_size_t] * 3),
            "orc_matrix_add": (None, [C.c_void_p, u64p, u64p, u64p, C.c_size_t]),
            "orc_matrix_sub": (None, [C.c_void_p, u64p, u64p, u64p, C.c_size_t]),
            "orc_matrix_mul_scalar": (None, [C.c_void_p, u64p, u64p, u64p, C.c_size_t]),
            "orc_crt_bits": (C.c_uint32, [C.c_void_p]),
            "orc_digits_per_tower": (C.c_uint32, [C.c_void_p, C.c_uint32]),
            "orc_matrix_decompose": (None, [C.c_void_p, u64p, u64p, C.c_size_t, C.c_size_t, C.c_uint32, C.c_int]),
            "orc_matrix_gadget": (None, [C.c_void_p, u64p, C.c_size_t, C.c_uint32, C.c_int]),
            "orc_matrix_mul_decompose": (None, [C.c_void_p, u64p, u64p, u64p] + [C.c_size_t] * 3 + [C.c_uint32, C.c_int]),
            "orc_chacha20_block": (None, [u32p, u32p, u32p]),
            "orc_hchacha20": (None, [u32p, u32p, u32p]),
            "orc_stream_words": (None, [u64p] + [C.c_uint64] * 4 + [C.c_size_t, u64p]),
            "orc_karney_samples": (None, [u64p, C.c_double, C.c_double, C.c_size_t, C.POINTER(C.c_int64)]),
            "orc_sample_distribution": (None, [C.c_int, C.c_double, u64p] + [C.c_size_t] * 5 + [C.c_uint32, u64p, u64p]),
            "orc_gauss_samp_gq": (None, [u64p] + [C.c_size_t] * 3 + [C.c_uint32, u64p, C.c_uint32, C.c_uint32, C.c_double, u64p, u64p]),
            "orc_sample_p1": (None, [u64p] * 4 + [C.c_size_t] * 3 + [C.c_uint32, u64p] + [C.c_double] * 3 + [u64p, u64p]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def prime_chain(n: int, bits: int, depth: int) -> list[int]:
    out = np.zeros(depth, dtype=np.uint64)
    k = lib().orc_prime_chain(n, bits, depth, _p(out))
    if k != depth:
        raise ValueError(f"only {k} of {depth} primes below 2^{bits} with p = 1 mod {2 * n}")
    return [int(x) for x in out]


def seed_words(seed) -> np.ndarray:
    """GpuRngSeed words: 4 ints, 32 LE bytes, or one int (word 0)."""
    if isinstance(seed, (bytes, bytearray)):
        seed = [int.from_bytes(seed[8 * i:8 * i + 8], "little") for i in range(4)]
    elif isinstance(seed, int):
        seed = [seed, 0, 0, 0]
    return np.array([int(w) & (2**64 - 1) for w in seed], dtype=np.uint64)


def chacha20_block(key_words, w12_15) -> np.ndarray:
    k, w, out = np.array(key_words, dtype=np.uint32), np.array(w12_15, dtype=np.uint32), np.zeros(16, dtype=np.uint32)
    lib().orc_chacha20_block(k.ctypes.data_as(u32p), w.ctypes.data_as(u32p), out.ctypes.data_as(u32p))
    return out


def hchacha20(key_words, nonce_words) -> np.ndarray:
    k, w, out = np.array(key_words, dtype=np.uint32), np.array(nonce_words, dtype=np.uint32), np.zeros(8, dtype=np.uint32)
    lib().orc_hchacha20(k.ctypes.data_as(u32p), w.ctypes.data_as(u32p), out.ctypes.data_as(u32p))
    return out


def stream_words(seed, stream0, stream1, stream2, domain, count) -> np.ndarray:
    out = np.zeros(count, dtype=np.uint64)
    lib().orc_stream_words(_p(seed_words(seed)), stream0, stream1, stream2, domain, count, _p(out))
    return out


def karney_samples(seed, mean, stddev, count) -> np.ndarray:
    out = np.zeros(count, dtype=np.int64)
    lib().orc_karney_samples(_p(seed_words(seed)), float(mean), float(stddev), count, out.ctypes.data_as(C.POINTER(C.c_int64)))
    return out


class Plan:
    """Per-(n, moduli) tables + the matrix-level oracle operations."""

    def __init__(self, n: int, moduli):
        self.n = int(n)
        self.moduli = [int(q) for q in moduli]
        self.L = len(self.moduli)
        self._mod = np.array(self.moduli, dtype=np.uint64)
        self._h = lib().orc_plan_create(self.n, _p(self._mod), self.L)
        if not self._h:
            raise ValueError("orc_plan_create failed (modulus not 1 mod 2n?)")
        self.crt_bits = int(lib().orc_crt_bits(self._h))

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.orc_plan_destroy(self._h)
            self._h = None

    # -- helpers ---------------------------------------------------------
    def dpt(self, base_bits: int) -> int:
        return -(-self.crt_bits // base_bits)

    def ell(self, base_bits: int, small: bool = False) -> int:
        return self.dpt(base_bits) * (1 if small else self.L)

    def uniform(self, rows: int, cols: int, seed: int) -> np.ndarray:
        m = np.empty((rows, cols, self.L, self.n), dtype=np.uint64)
        lib().orc_fill_uniform(_p(m), rows * cols, self.L, self.n, _p(self._mod), seed)
        return m

    def _chk(self, m: np.ndarray) -> np.ndarray:
        m = np.ascontiguousarray(m, dtype=np.uint64)
        assert m.ndim == 4 and m.shape[2] == self.L and m.shape[3] == self.n, m.shape
        return m

    # -- samplers (sampler_oracle.c), all coefficient domain ----------------
    def sample_distribution(self, rows, cols, dist, sigma, seed, full_cols=None, col_offset=0) -> np.ndarray:
        out = np.empty((rows, cols, self.L, self.n), dtype=np.uint64)
        lib().orc_sample_distribution(dist, float(sigma), _p(seed_words(seed)), rows, cols, cols if full_cols is None else full_cols,
                                      col_offset, self.n, self.L, _p(self._mod), _p(out))
        return out

    def gauss_samp_gq(self, src_coeff: np.ndarray, base_bits: int, c: float, seed) -> np.ndarray:
        src = self._chk(src_coeff)
        rows, cols = src.shape[:2]
        out = np.empty((rows * self.ell(base_bits), cols, self.L, self.n), dtype=np.uint64)
        lib().orc_gauss_samp_gq(_p(src), rows, cols, self.n, self.L, _p(self._mod), self.crt_bits, base_bits, float(c),
                                _p(seed_words(seed)), _p(out))
        return out

    def sample_p1(self, a, b, dm, tp2, sigma, s, dgg_stddev, seed) -> np.ndarray:
        a, b, dm, tp2 = (self._chk(x) for x in (a, b, dm, tp2))
        d, cols = a.shape[0], tp2.shape[1]
        assert tp2.shape[0] == 2 * d
        out = np.empty_like(tp2)
        lib().orc_sample_p1(_p(a), _p(b), _p(dm), _p(tp2), d, cols, self.n, self.L, _p(self._mod), float(sigma), float(s),
                            float(dgg_stddev), _p(seed_words(seed)), _p(out))
        return out

    # -- transforms ------------------------------------------------------
    def ntt(self, m: np.ndarray) -> np.ndarray:
        out = self._chk(m).copy()
        lib().orc_matrix_ntt(self._h, _p(out), out.shape[0] * out.shape[1])
        return out

    def intt(self, m: np.ndarray) -> np.ndarray:
        out = self._chk(m).copy()
        lib().orc_matrix_intt(self._h, _p(out), out.shape[0] * out.shape[1])
        return out

    # -- arithmetic ------------------------------------------------------
    def mul(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        a, b = self._chk(a), self._chk(b)
        r, k = a.shape[:2]
        k2, c = b.shape[:2]
        assert k == k2
        out = np.empty((r, c, self.L, self.n), dtype=np.uint64)
        lib().orc_matrix_mul(self._h, _p(out), _p(a), _p(b), r, k, c)
        return out

    def add(self, a, b):
        a, b = self._chk(a), self._chk(b)
        out = np.empty_like(a)
        lib().orc_matrix_add(self._h, _p(out), _p(a), _p(b), a.shape[0] * a.shape[1])
        return out

    def sub(self, a, b):
        a, b = self._chk(a), self._chk(b)
        out = np.empty_like(a)
        lib().orc_matrix_sub(self._h, _p(out), _p(a), _p(b), a.shape[0] * a.shape[1])
        return out

    def mul_scalar(self, a, s):
        a, s = self._chk(a), self._chk(s)
        assert s.shape[:2] == (1, 1)
        out = np.empty_like(a)
        lib().orc_matrix_mul_scalar(self._h, _p(out), _p(a), _p(s), a.shape[0] * a.shape[1])
        return out

    # -- gadget ----------------------------------------------------------
    def decompose(self, src_coeff: np.ndarray, base_bits: int, small: bool = False) -> np.ndarray:
        src = self._chk(src_coeff)
        rows, cols = src.shape[:2]
        out = np.empty((rows * self.ell(base_bits, small), cols, self.L, self.n), dtype=np.uint64)
        lib().orc_matrix_decompose(self._h, _p(out), _p(src), rows, cols, base_bits, int(small))
        return out

    def gadget(self, rows: int, base_bits: int, small: bool = False) -> np.ndarray:
        out = np.empty((rows, rows * self.ell(base_bits, small), self.L, self.n), dtype=np.uint64)
        lib().orc_matrix_gadget(self._h, _p(out), rows, base_bits, int(small))
        return out

    def mul_decompose(self, S: np.ndarray, other: np.ndarray, base_bits: int, small: bool = False) -> np.ndarray:
        S, other = self._chk(S), self._chk(other)
        r, m = S.shape[:2]
        d, c = other.shape[:2]
        assert m == d * self.ell(base_bits, small), (m, d, self.ell(base_bits, small))
        out = np.empty((r, c, self.L, self.n), dtype=np.uint64)
        lib().orc_matrix_mul_decompose(self._h, _p(out), _p(S), _p(other), r, d, c, base_bits, int(small))
        return out


# -- single-limb reference transforms (slow; small n) -------------------------
def ntt_naive(a: np.ndarray, q: int, psi: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    out = np.empty_like(a)
    lib().orc_ntt_naive(_p(a), _p(out), a.size, q, psi)
    return out


def ntt_ref(a: np.ndarray, q: int, psi: int) -> np.ndarray:
    out = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_ntt_ref(_p(out), out.size, q, psi)
    return out


def intt_ref(a: np.ndarray, q: int, psi: int) -> np.ndarray:
    out = np.ascontiguousarray(a, dtype=np.uint64).copy()
    lib().orc_intt_ref(_p(out), out.size, q, psi)
    return out


def negacyclic_schoolbook(a: np.ndarray, b: np.ndarray, q: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = np.ascontiguousarray(b, dtype=np.uint64)
    c = np.empty_like(a)
    lib().orc_negacyclic_schoolbook(_p(a), _p(b), _p(c), a.size, q)
    return c


# -- compact wire format (pure-Python big integers; small cases only) --------------------------
def compact_encode(coeff_rns: np.ndarray, moduli) -> tuple[int, bytes]:
    """Restates gpu_poly_store_compact_bytes (cuda/src/matrix/MatrixSerde.cu:1066-1630) on a COEFFICIENT-form
    matrix (nrow, ncol, limbs, n): CRT-reconstruct every coefficient to X in [0, Q) (kernel :162-278),
    centre it — negative iff X > floor(Q/2), magnitude Q - X (:280-338) — and pack B = max magnitude width + 1
    bits per coefficient, LSB first, the sign in bit B-1 (:340-386); B = 0 when everything is zero (:1552).
    Returns (B, payload)."""
    m = np.ascontiguousarray(coeff_rns, dtype=np.uint64)
    nrow, ncol, L, n = m.shape
    qs = [int(q) for q in moduli[:L]]
    Q = 1
    for q in qs:
        Q *= q
    basis = [(Q // q) * pow(Q // q, -1, q) for q in qs]
    half = Q >> 1
    vals = []
    flat = m.reshape(nrow * ncol, L, n)
    for p in range(flat.shape[0]):
        for j in range(n):
            X = sum(int(flat[p, l, j]) * basis[l] for l in range(L)) % Q
            vals.append((1, Q - X) if X > half else (0, X))
    width = max((mag.bit_length() for _, mag in vals), default=0)
    if width == 0:
        return 0, b""
    B = width + 1
    acc = 0
    for k, (neg, mag) in enumerate(vals):
        acc |= (mag | (neg << (B - 1))) << (k * B)
    return B, acc.to_bytes((len(vals) * B + 7) // 8, "little")


def compact_decode(payload: bytes, B: int, nrow: int, ncol: int, moduli, n: int) -> np.ndarray:
    """Restates gpu_poly_load_compact_bytes / serde_unpack_packed_coeffs_mod_kernel (MatrixSerde.cu:388-470, 1632-1944):
    residues of the signed coefficient in every limb, coefficient form."""
    L = len(moduli)
    out = np.zeros((nrow * ncol, L, n), dtype=np.uint64)
    if B:
        acc = int.from_bytes(payload, "little")
        mask = (1 << (B - 1)) - 1
        for k in range(nrow * ncol * n):
            v = (acc >> (k * B)) & ((1 << B) - 1)
            mag, neg = v & mask, v >> (B - 1)
            for l, q in enumerate(moduli):
                r = mag % int(q)
                out[k // n, l, k % n] = (int(q) - r) % int(q) if neg else r
    return out.reshape(nrow, ncol, L, n)
