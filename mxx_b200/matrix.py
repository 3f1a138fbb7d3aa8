"""GpuDCRTPolyMatrix — host mirror of src/matrix/gpu_dcrt_poly.rs (the `PolyMatrix`
surface of src/matrix/mod.rs:45-379 over the gpu_* C ABI).

Same method names, argument meaning and failure behaviour as the Rust type so the parity
tests read like the reference's own (`gpu_dcrt_poly.rs:1926-2700`).  Host-side matrices
are numpy uint64 arrays of shape (nrow, ncol, level+1, n): the flat
[poly row-major][limb][coeff] little-endian u64 layout of gpu_dcrt_poly.rs:769-817,
1916-1920.  All compute happens in libgpupoly.so; nothing here touches residues.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import ffi
from .ffi import GPU_POLY_FORMAT_COEFF, GPU_POLY_FORMAT_EVAL, GpuRngSeed, check_status
from .params import GpuDCRTPolyParams


def mul_decompose_column_chunk_width() -> int:
    """src/env.rs:148-154: MXX_MUL_DECOMPOSE_COLUMN_CHUNK_WIDTH, default 1."""
    try:
        return max(1, int(os.environ.get("MXX_MUL_DECOMPOSE_COLUMN_CHUNK_WIDTH", "1")))
    except ValueError:
        return 1



def _varint(v: int) -> bytes:
    """bincode 2 standard config: unsigned varint (u16/u32/u64/usize)"""
    v = int(v)
    if v < 251:
        return bytes([v])
    if v < 1 << 16:
        return bytes([251]) + v.to_bytes(2, "little")
    if v < 1 << 32:
        return bytes([252]) + v.to_bytes(4, "little")
    return bytes([253]) + v.to_bytes(8, "little")


def _unvarint(data: bytes, pos: int):
    t = data[pos]
    if t < 251:
        return t, pos + 1
    width = {251: 2, 252: 4, 253: 8}[t]
    return int.from_bytes(data[pos + 1: pos + 1 + width], "little"), pos + 1 + width


class PinnedBuffer:
    """Page-locked host buffer from gpu_pinned_alloc (gpu.rs:236-237), viewed as numpy uint64."""

    def __init__(self, shape, library=None):
        self._L = library or ffi.lib()
        self.shape = tuple(int(s) for s in shape)
        self.nbytes = int(np.prod(self.shape)) * 8
        self._ptr = self._L.gpu_pinned_alloc(self.nbytes) if self.nbytes else None
        if self.nbytes and not self._ptr:
            check_status(1, "gpu_pinned_alloc", self._L)
        buf = (C.c_uint64 * (self.nbytes // 8)).from_address(self._ptr) if self.nbytes else (C.c_uint64 * 0)()
        self.array = np.frombuffer(buf, dtype=np.uint64).reshape(self.shape)

    def free(self):
        if getattr(self, "_ptr", None):
            self.array = None
            self._L.gpu_pinned_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class GpuDCRTPolyMatrix:
    def __init__(self, params: GpuDCRTPolyParams, nrow: int, ncol: int, level: int | None = None, is_ntt: bool = True):
        """new_empty_with_state (gpu_dcrt_poly.rs:222-256): uninitialised device storage."""
        self.params = params
        self._L = params._L
        self.nrow, self.ncol = int(nrow), int(ncol)
        self.level = params.crt_depth - 1 if level is None else int(level)
        assert self.level < params.crt_depth, "invalid level for matrix create"
        self.is_ntt = bool(is_ntt)
        h = C.c_void_p()
        fmt = GPU_POLY_FORMAT_EVAL if is_ntt else GPU_POLY_FORMAT_COEFF
        check_status(self._L.gpu_matrix_create(params.ctx_raw(), self.level, self.nrow, self.ncol, fmt, C.byref(h)),
                     "gpu_matrix_create", self._L)
        self.raw = h

    # -- lifetime ---------------------------------------------------------------------
    def destroy(self):
        if getattr(self, "raw", None):
            self._L.gpu_matrix_destroy(self.raw)
            self.raw = None

    def __del__(self):  # Drop, gpu_dcrt_poly.rs:145-152
        try:
            self.destroy()
        except Exception:
            pass

    # -- constructors -------------------------------------------------------------------
    @classmethod
    def new_empty(cls, params, nrow, ncol):
        return cls(params, nrow, ncol)

    @classmethod
    def zero(cls, params, nrow, ncol, level=None, is_ntt=True):
        """new_zero_with_state (gpu_dcrt_poly.rs:344-365): load an all-zero host buffer."""
        out = cls(params, nrow, ncol, level, is_ntt)
        if nrow and ncol:
            host = np.zeros((nrow, ncol, out.level + 1, params.ring_dimension), dtype=np.uint64)
            out.load_rns(host, out._fmt())
        return out

    @classmethod
    def from_rns(cls, params, host: np.ndarray, is_ntt: bool = True):
        """from_cpu_matrix / from_rns_snapshot (gpu_dcrt_poly.rs:769-817, 681-698): host residues -> device."""
        host = np.ascontiguousarray(host, dtype=np.uint64)
        assert host.ndim == 4 and host.shape[3] == params.ring_dimension
        out = cls(params, host.shape[0], host.shape[1], host.shape[2] - 1, is_ntt)
        out.load_rns(host, out._fmt())
        return out

    @classmethod
    def identity(cls, params, size, scalar: "GpuDCRTPolyMatrix | None" = None):
        """gpu_dcrt_poly.rs:1158-1188: size x size, `scalar` (a 1x1 matrix, default 1) on the diagonal."""
        n, L = params.ring_dimension, params.crt_depth
        host = np.zeros((size, size, L, n), dtype=np.uint64)
        if scalar is None:
            diag = np.ones((L, n), dtype=np.uint64)  # the constant 1 evaluates to 1 at every slot
        else:
            s = scalar.clone()
            s.ntt_in_place()
            diag = s.to_rns()[0, 0]
        for i in range(size):
            host[i, i] = diag
        return cls.from_rns(params, host, True)

    @classmethod
    def gadget_matrix(cls, params, size):
        """gpu_dcrt_poly.rs:1254-1263"""
        if size == 0:
            return cls(params, 0, 0)
        out = cls(params, size, size * params.modulus_digits())
        check_status(out._L.gpu_matrix_fill_gadget(out.raw, params.base_bits), "gpu_matrix_fill_gadget", out._L)
        out.is_ntt = True
        return out

    @classmethod
    def small_gadget_matrix(cls, params, size):
        """gpu_dcrt_poly.rs:1265-1274"""
        if size == 0:
            return cls(params, 0, 0)
        out = cls(params, size, size * params.digits_per_tower)
        check_status(out._L.gpu_matrix_fill_small_gadget(out.raw, params.base_bits), "gpu_matrix_fill_small_gadget", out._L)
        out.is_ntt = True
        return out

    @classmethod
    def small_decomposed_identity_chunk(cls, params, size, chunk_idx, chunk_count, scalar_by_digit: "GpuDCRTPolyMatrix"):
        """gpu_dcrt_poly.rs:1296-1326; scalar_by_digit is the 1 x chunk_count row of digit polynomials."""
        assert chunk_count > 0 and scalar_by_digit.size() == (1, chunk_count)
        if size == 0:
            return cls(params, 0, 0)
        out = cls(params, size, size, scalar_by_digit.level, scalar_by_digit.is_ntt)
        check_status(out._L.gpu_matrix_fill_small_decomposed_identity_chunk(out.raw, scalar_by_digit.raw, chunk_idx),
                     "gpu_matrix_fill_small_decomposed_identity_chunk", out._L)
        return out

    # -- host <-> device ------------------------------------------------------------------
    def _fmt(self):
        return GPU_POLY_FORMAT_EVAL if self.is_ntt else GPU_POLY_FORMAT_COEFF

    def _wait(self, events, what):
        if events:  # gpu_dcrt_poly.rs:591-595
            st = self._L.gpu_event_set_wait(events)
            self._L.gpu_event_set_destroy(events)
            check_status(st, "gpu_event_set_wait", self._L)

    def bytes_per_poly(self) -> int:
        return (self.level + 1) * self.params.ring_dimension * 8  # gpu_dcrt_poly.rs:1916-1920

    def load_rns(self, host: np.ndarray, fmt: int, wait: bool = True):
        """load_rns_bytes (gpu_dcrt_poly.rs:629-663)."""
        assert host.dtype == np.uint64 and host.flags["C_CONTIGUOUS"]
        if host.size == 0:
            return None
        assert host.size * 8 == self.nrow * self.ncol * self.bytes_per_poly()
        ev = C.c_void_p()
        check_status(self._L.gpu_matrix_load_rns_batch(self.raw, host.ctypes.data, self.bytes_per_poly(), fmt, C.byref(ev)),
                     "gpu_matrix_load_rns_batch", self._L)
        self.is_ntt = fmt == GPU_POLY_FORMAT_EVAL
        if wait:
            self._wait(ev, "load")
            return None
        return ev

    def store_rns(self, out: np.ndarray, fmt: int | None = None, wait: bool = True):
        """store_rns_bytes (gpu_dcrt_poly.rs:576-596): the tag must equal the current domain."""
        assert out.dtype == np.uint64 and out.flags["C_CONTIGUOUS"]
        if out.size == 0:
            return None
        ev = C.c_void_p()
        check_status(self._L.gpu_matrix_store_rns_batch(self.raw, out.ctypes.data, self.bytes_per_poly(),
                                                        self._fmt() if fmt is None else fmt, C.byref(ev)),
                     "gpu_matrix_store_rns_batch", self._L)
        if wait:
            self._wait(ev, "store")
            return None
        return ev

    def to_rns(self) -> np.ndarray:
        """to_rns_snapshot (gpu_dcrt_poly.rs:665-679): residues in the matrix's current domain."""
        out = np.empty((self.nrow, self.ncol, self.level + 1, self.params.ring_dimension), dtype=np.uint64)
        self.store_rns(out)
        return out

    def store_const_coeff_words(self) -> np.ndarray:
        """gpu_dcrt_poly.rs:598-627: coefficient 0 of every limb; requires the coefficient domain."""
        words = np.zeros((self.nrow, self.ncol, self.level + 1), dtype=np.uint64)
        if words.size == 0:
            return words
        ev = C.c_void_p()
        check_status(self._L.gpu_matrix_store_const_coeff_batch(self.raw, words.ctypes.data_as(ffi.u64p), self.level + 1, C.byref(ev)),
                     "gpu_matrix_store_const_coeff_batch", self._L)
        self._wait(ev, "store_const")
        return words

    # -- big-integer coefficients (device-side conversion, extension) ------------------------
    @classmethod
    def from_biguints(cls, params, values, nrow: int, ncol: int, is_ntt: bool = True, level: int | None = None):
        """from_cpu_matrix / from_coeffs (gpu_dcrt_poly.rs:769-817, gpu.rs:903-920) with the `value % modulus`
        loop done on the device: `values` = nrow*ncol*n Python ints (row-major polys, natural coefficient /
        slot order), any non-negative size."""
        n = params.ring_dimension
        vals = [int(v) for v in values]
        assert len(vals) == nrow * ncol * n and all(v >= 0 for v in vals)
        wpc = max(1, (max((v.bit_length() for v in vals), default=1) + 63) // 64)
        words = np.zeros((len(vals), wpc), dtype=np.uint64)
        mask = (1 << 64) - 1
        for i, v in enumerate(vals):
            for w in range(wpc):
                words[i, w] = (v >> (64 * w)) & mask
        out = cls(params, nrow, ncol, level, is_ntt)
        ev = C.c_void_p()
        check_status(out._L.gpu_matrix_load_bigint_batch(out.raw, words.ctypes.data_as(ffi.u64p), wpc, out._fmt(), C.byref(ev)),
                     "gpu_matrix_load_bigint_batch", out._L)
        out._wait(ev, "load_bigint")
        return out

    def to_biguints(self) -> list:
        """coeffs() (gpu.rs:959-994) for every polynomial: CRT-reconstructed coefficients in [0, Q) as Python ints,
        reconstructed on the device (a coefficient-form copy is made if the matrix is in the evaluation domain)."""
        src = self
        if self.is_ntt:
            src = self.clone()
            src.intt_all_in_place()
        n = self.params.ring_dimension
        q_bits = sum(q.bit_length() for q in self.params.moduli[: self.level + 1])
        wpc = max(1, (q_bits + 63) // 64)
        count = self.nrow * self.ncol * n
        words = np.zeros((count, wpc), dtype=np.uint64)
        if count == 0:
            return []
        ev = C.c_void_p()
        check_status(self._L.gpu_matrix_store_bigint_batch(src.raw, words.ctypes.data_as(ffi.u64p), wpc, C.byref(ev)),
                     "gpu_matrix_store_bigint_batch", self._L)
        self._wait(ev, "store_bigint")
        return [sum(int(words[i, w]) << (64 * w) for w in range(wpc)) for i in range(count)]

    # -- compact wire format (gpu_dcrt_poly.rs:956-1044, 1681-1712) -----------------------------------
    def to_compact_bytes(self) -> bytes:
        """into_compact_bytes on a clone (the native call converts an EVAL matrix to coefficient form in
        place, MatrixSerde.cu:1108-1115; the Rust method consumes `self`).  Envelope: bincode standard config of
        (1u8, format u8, level u32, nrow, ncol, max_coeff_bits u16, bytes_per_coeff u16, payload)."""
        src = self.clone()
        n = self.params.ring_dimension
        count = self.nrow * self.ncol * n
        bits_upper = sum(q.bit_length() for q in self.params.moduli[: self.level + 1])
        capacity = (count * bits_upper + 7) // 8
        buf = (C.c_uint8 * max(capacity, 1))()
        mb, bpc, plen = C.c_uint16(0), C.c_uint16(0), C.c_size_t(0)
        check_status(self._L.gpu_matrix_store_compact_bytes(src.raw, buf, capacity, C.byref(mb), C.byref(bpc), C.byref(plen)),
                     "gpu_matrix_store_compact_bytes", self._L)
        payload = bytes(buf[: plen.value])
        fmt = GPU_POLY_FORMAT_EVAL if self.is_ntt else GPU_POLY_FORMAT_COEFF
        return (bytes([1, fmt]) + _varint(self.level) + _varint(self.nrow) + _varint(self.ncol) + _varint(mb.value) + _varint(bpc.value)
                + _varint(len(payload)) + payload)

    @classmethod
    def from_compact_bytes(cls, params, data: bytes):
        """gpu_dcrt_poly.rs:1004-1044"""
        version, fmt = data[0], data[1]
        assert version == 1, f"Unsupported compact matrix version: {version}"
        assert fmt in (GPU_POLY_FORMAT_COEFF, GPU_POLY_FORMAT_EVAL), f"Invalid compact matrix format tag: {fmt}"
        pos = 2
        level, pos = _unvarint(data, pos)
        nrow, pos = _unvarint(data, pos)
        ncol, pos = _unvarint(data, pos)
        max_bits, pos = _unvarint(data, pos)
        bpc, pos = _unvarint(data, pos)
        plen, pos = _unvarint(data, pos)
        payload = data[pos: pos + plen]
        assert len(payload) == plen, "truncated compact payload"
        assert level < params.crt_depth, f"invalid compact matrix level: {level}"
        assert bpc == (max_bits + 7) // 8, f"compact bytes_per_coeff mismatch: got {bpc}, expected {(max_bits + 7) // 8}"
        out = cls(params, nrow, ncol, level, False)
        buf = (C.c_uint8 * max(plen, 1)).from_buffer_copy(payload if plen else b"\0")
        check_status(out._L.gpu_matrix_load_compact_bytes(out.raw, buf, plen, max_bits), "gpu_matrix_load_compact_bytes", out._L)
        out.is_ntt = False
        if fmt == GPU_POLY_FORMAT_EVAL:
            out.ntt_in_place()
        return out

    @staticmethod
    def zero_compact_bytes(params, nrow, ncol, level, is_ntt, max_coeff_bits) -> bytes:
        """gpu_dcrt_poly.rs:1681-1712"""
        assert level < params.crt_depth, "invalid level for compact zero matrix"
        mb = max(int(max_coeff_bits), 1)
        plen = (nrow * ncol * params.ring_dimension * mb + 7) // 8
        fmt = GPU_POLY_FORMAT_EVAL if is_ntt else GPU_POLY_FORMAT_COEFF
        return bytes([1, fmt]) + _varint(level) + _varint(nrow) + _varint(ncol) + _varint(mb) + _varint((mb + 7) // 8) + _varint(plen) + bytes(plen)

    # -- shape / domain --------------------------------------------------------------------
    def size(self):
        return (self.nrow, self.ncol)

    def clone(self):
        out = GpuDCRTPolyMatrix(self.params, self.nrow, self.ncol, self.level, self.is_ntt)
        check_status(self._L.gpu_matrix_copy(out.raw, self.raw), "gpu_matrix_copy", self._L)
        return out

    def ntt_in_place(self):
        check_status(self._L.gpu_matrix_ntt_all(self.raw), "gpu_matrix_ntt_all", self._L)
        self.is_ntt = True

    def intt_all_in_place(self):  # gpu_dcrt_poly.rs:295-302
        check_status(self._L.gpu_matrix_intt_all(self.raw), "gpu_matrix_intt_all", self._L)
        self.is_ntt = False

    def into_coeff_domain(self):
        self.intt_all_in_place()
        return self

    def __eq__(self, other):  # PartialEq, gpu_dcrt_poly.rs:199-217
        if not isinstance(other, GpuDCRTPolyMatrix):
            return NotImplemented
        if (self.params is not other.params and self.params.moduli != other.params.moduli) or self.size() != other.size() \
                or self.level != other.level or self.is_ntt != other.is_ntt:
            return False
        if self.raw.value == other.raw.value:
            return True
        eq = C.c_int(0)
        check_status(self._L.gpu_matrix_equal(self.raw, other.raw, C.byref(eq)), "gpu_matrix_equal", self._L)
        return eq.value != 0

    __hash__ = None

    # -- blocks -------------------------------------------------------------------------------
    def copy_block_from(self, src, dst_row, dst_col, src_row, src_col, rows, cols):  # gpu_dcrt_poly.rs:372-389
        check_status(self._L.gpu_matrix_copy_block(self.raw, src.raw, dst_row, dst_col, src_row, src_col, rows, cols),
                     "gpu_matrix_copy_block", self._L)
        self.is_ntt = src.is_ntt

    def add_block_from(self, src, dst_row, dst_col, src_row, src_col, rows, cols):  # gpu_dcrt_poly.rs:411-432
        assert self.is_ntt == src.is_ntt, "add_block_from requires matching domains"
        check_status(self._L.gpu_matrix_add_block(self.raw, src.raw, dst_row, dst_col, src_row, src_col, rows, cols),
                     "gpu_matrix_add_block", self._L)

    def slice(self, row_start, row_end, col_start, col_end):  # gpu_dcrt_poly.rs:1146-1152
        out = GpuDCRTPolyMatrix(self.params, row_end - row_start, col_end - col_start, self.level, self.is_ntt)
        out.copy_block_from(self, 0, 0, row_start, col_start, row_end - row_start, col_end - col_start)
        return out

    def slice_rows(self, start, end):
        return self.slice(start, end, 0, self.ncol)

    def slice_columns(self, start, end):
        return self.slice(0, self.nrow, start, end)

    def entry(self, i, j):
        """1x1 matrix holding entry (i, j) — the GpuDCRTPoly of gpu.rs:707 is a 1x1 matrix wrapper."""
        return self.slice(i, i + 1, j, j + 1)

    def set_entry(self, i, j, elem):
        assert elem.size() == (1, 1)
        e = elem
        if e.is_ntt != self.is_ntt:
            e = elem.clone()
            e.ntt_in_place() if self.is_ntt else e.intt_all_in_place()
        self.copy_block_from(e, i, j, 0, 0, 1, 1)

    def transpose(self):  # gpu_dcrt_poly.rs:1190-1199
        out = GpuDCRTPolyMatrix(self.params, self.ncol, self.nrow, self.level, self.is_ntt)
        for i in range(self.nrow):
            for j in range(self.ncol):
                out.copy_block_from(self, j, i, i, j, 1, 1)
        return out

    def _same_domain(self, others):
        for o in others:
            assert o.level == self.level and o.is_ntt == self.is_ntt, "concat requires same level and domain"

    def concat_columns(self, others):  # gpu_dcrt_poly.rs:853-885
        self._same_domain(others)
        assert all(o.nrow == self.nrow for o in others)
        out = GpuDCRTPolyMatrix(self.params, self.nrow, self.ncol + sum(o.ncol for o in others), self.level, self.is_ntt)
        c = 0
        for m in [self, *others]:
            out.copy_block_from(m, 0, c, 0, 0, m.nrow, m.ncol)
            c += m.ncol
        return out

    def concat_rows(self, others):  # gpu_dcrt_poly.rs:819-851
        self._same_domain(others)
        assert all(o.ncol == self.ncol for o in others)
        out = GpuDCRTPolyMatrix(self.params, self.nrow + sum(o.nrow for o in others), self.ncol, self.level, self.is_ntt)
        r = 0
        for m in [self, *others]:
            out.copy_block_from(m, r, 0, 0, 0, m.nrow, m.ncol)
            r += m.nrow
        return out

    def concat_diag(self, others):  # gpu_dcrt_poly.rs:887-916
        self._same_domain(others)
        mats = [self, *others]
        out = GpuDCRTPolyMatrix.zero(self.params, sum(m.nrow for m in mats), sum(m.ncol for m in mats), self.level, self.is_ntt)
        r = c = 0
        for m in mats:
            out.copy_block_from(m, r, c, 0, 0, m.nrow, m.ncol)
            r += m.nrow
            c += m.ncol
        return out

    # -- ring operations --------------------------------------------------------------------------
    def _binary(self, rhs, fn, name):
        assert self.size() == rhs.size(), f"{name} requires equal shapes"
        assert self.is_ntt == rhs.is_ntt, f"{name} requires matching domains"
        out = GpuDCRTPolyMatrix(self.params, self.nrow, self.ncol, self.level, self.is_ntt)
        check_status(fn(out.raw, self.raw, rhs.raw), name, self._L)
        return out

    def __add__(self, rhs):  # gpu_dcrt_poly.rs:1716-1741
        return self._binary(rhs, self._L.gpu_matrix_add, "gpu_matrix_add")

    def __sub__(self, rhs):  # gpu_dcrt_poly.rs:1744-1768
        return self._binary(rhs, self._L.gpu_matrix_sub, "gpu_matrix_sub")

    def __neg__(self):  # gpu_dcrt_poly.rs:1885-1899: zero - self
        return GpuDCRTPolyMatrix.zero(self.params, self.nrow, self.ncol, self.level, self.is_ntt) - self

    def add_in_place(self, rhs):  # gpu_dcrt_poly.rs:391-409
        assert self.size() == rhs.size() and self.is_ntt == rhs.is_ntt
        check_status(self._L.gpu_matrix_add(self.raw, self.raw, rhs.raw), "gpu_matrix_add", self._L)

    def _eval(self):
        """operand in the evaluation domain (a converted copy if needed; the original is not modified)"""
        if self.is_ntt:
            return self
        c = self.clone()
        c.ntt_in_place()
        return c

    def mul_scalar(self, scalar):  # gpu_dcrt_poly.rs:1770-1790 (scalar: a 1x1 matrix)
        assert scalar.size() == (1, 1)
        lhs, s = self._eval(), scalar._eval()
        out = GpuDCRTPolyMatrix(self.params, self.nrow, self.ncol, self.level, True)
        check_status(self._L.gpu_matrix_mul_scalar(out.raw, lhs.raw, s.raw), "gpu_matrix_mul_scalar", self._L)
        return out

    def __mul__(self, rhs):  # mul_internal, gpu_dcrt_poly.rs:1792-1815; Mul<P> :1853-1883
        if rhs.size() == (1, 1) and self.ncol != 1:
            return self.mul_scalar(rhs)
        assert self.ncol == rhs.nrow, "matrix product requires lhs.ncol == rhs.nrow"
        lhs, r = self._eval(), rhs._eval()
        out = GpuDCRTPolyMatrix(self.params, self.nrow, rhs.ncol, self.level, True)
        check_status(self._L.gpu_matrix_mul(out.raw, lhs.raw, r.raw), "gpu_matrix_mul", self._L)
        return out

    __matmul__ = __mul__

    def tensor(self, other):  # gpu_dcrt_poly.rs:1225-1252
        out = GpuDCRTPolyMatrix(self.params, self.nrow * other.nrow, self.ncol * other.ncol, self.level, True)
        for i in range(self.nrow):
            for j in range(self.ncol):
                block = other.mul_scalar(self.entry(i, j))
                out.copy_block_from(block, i * other.nrow, j * other.ncol, 0, 0, other.nrow, other.ncol)
        return out

    # -- gadget decomposition ------------------------------------------------------------------------
    def _decompose(self, small: bool):
        """decompose_from_raw (gpu_dcrt_poly.rs:309-328): the result comes back in the evaluation domain."""
        p = self.params
        ell = p.digits_per_tower if small else p.digits_per_tower * (self.level + 1)
        out = GpuDCRTPolyMatrix(p, self.nrow * ell, self.ncol, self.level, True)
        fn = self._L.gpu_matrix_decompose_base_small if small else self._L.gpu_matrix_decompose_base
        check_status(fn(self.raw, p.base_bits, out.raw), "gpu_matrix_decompose_base", self._L)
        return out

    def decompose_noalloc(self, out=None):
        """gpu_matrix_decompose_base_noalloc (no counterpart in the reference): the same decomposition with the scratch in a
        static device array; `out` may be a matrix created earlier so that the call itself allocates nothing at all."""
        p = self.params
        ell = p.digits_per_tower * (self.level + 1)
        if out is None:
            out = GpuDCRTPolyMatrix(p, self.nrow * ell, self.ncol, self.level, True)
        check_status(self._L.gpu_matrix_decompose_base_noalloc(self.raw, p.base_bits, out.raw), "gpu_matrix_decompose_base_noalloc", self._L)
        return out

    def decompose(self):  # gpu_dcrt_poly.rs:1276-1284
        return self._decompose(False)

    def small_decompose(self):  # gpu_dcrt_poly.rs:1286-1294
        return self._decompose(True)

    def get_column_matrix_decompose(self, j):  # gpu_dcrt_poly.rs:1576-1579
        return self.slice(0, self.nrow, j, j + 1).decompose()

    def mul_decompose(self, other, fused: bool | None = None, small: bool = False):
        """S * G^-1(other) (gpu_dcrt_poly.rs:1414-1493; small: :1495-1574).

        fused=False replays the reference's call sequence verbatim (column chunks of
        MXX_MUL_DECOMPOSE_COLUMN_CHUNK_WIDTH: slice -> decompose -> mul -> copy_block), which is
        what the unchanged Rust side issues.  fused=True (default when the library exports it) is the
        single-call extension gpu_matrix_mul_decompose a maintainer would switch that loop to.
        """
        p = self.params
        ell = p.digits_per_tower if small else p.modulus_digits()
        assert self.ncol == other.nrow * ell, "mul_decompose requires self.ncol == other.nrow * log_base_q"
        ncol = other.ncol
        if self.nrow == 0 or ncol == 0:
            return GpuDCRTPolyMatrix(p, self.nrow, ncol)
        if fused is None:
            fused = hasattr(self._L, "gpu_matrix_mul_decompose")
        lhs = self._eval()
        out = GpuDCRTPolyMatrix(p, self.nrow, ncol, self.level, True)
        if fused:
            check_status(self._L.gpu_matrix_mul_decompose(out.raw, lhs.raw, other.raw, p.base_bits, int(small)),
                         "gpu_matrix_mul_decompose", self._L)
            return out
        width = max(1, min(mul_decompose_column_chunk_width(), ncol))
        for c0 in range(0, ncol, width):
            c1 = min(c0 + width, ncol)
            col = other.slice(0, other.nrow, c0, c1)
            dec = col._decompose(small)
            prod = lhs * dec
            out.copy_block_from(prod, 0, c0, 0, 0, self.nrow, c1 - c0)
        return out

    def mul_decompose_gathered(self, gathered, d: int, tower_moduli, towers_per_group: int, small: bool = False,
                               col0: int = 0, ncol: int | None = None):
        """S * G^-1(other) for a LIMB-SHARDED job (extension gpu_matrix_mul_decompose_gathered; mxx_b200/dist.py).

        self lives in this rank's context (its subset of the CRT basis).  `gathered` is the coefficient-form right
        operand with ALL towers, as the all-gather of the ranks' (d x c) blocks leaves it: a (groups*d) x c COEFF matrix
        of this context whose rows [g*d, (g+1)*d) are limb group g's towers.  Digit rows follow the full basis
        `tower_moduli`.  Returns the product on this rank's limbs, columns [col0, col0 + ncol) of `other`."""
        import ctypes as C

        towers = len(tower_moduli)
        assert towers % towers_per_group == 0 and gathered.nrow == (towers // towers_per_group) * d and not gathered.is_ntt
        assert gathered.level + 1 == towers_per_group == self.level + 1
        c = gathered.ncol
        ncol = c - col0 if ncol is None else ncol
        n = self.params.ring_dimension
        ptr, wb, fmt = C.c_void_p(), C.c_int(), C.c_int()
        check_status(self._L.gpu_matrix_device_info(gathered.raw, C.byref(ptr), C.byref(wb), C.byref(fmt)), "gpu_matrix_device_info", self._L)
        out = GpuDCRTPolyMatrix(self.params, self.nrow, ncol, self.level, True)
        mod = (C.c_uint64 * towers)(*[int(q) for q in tower_moduli])
        poly_stride = towers_per_group * n
        check_status(self._L.gpu_matrix_mul_decompose_gathered(out.raw, self._eval().raw, ptr, d, c, col0, mod, towers, towers_per_group,
                                                               d * c * poly_stride, poly_stride, self.params.base_bits, int(small)),
                     "gpu_matrix_mul_decompose_gathered", self._L)
        return out

    def intt_rows_from(self, src, dst_row: int = 0):
        """rows [dst_row, dst_row + src.nrow) of self <- coefficient form of src, src untouched (extension gpu_matrix_intt_to)"""
        check_status(self._L.gpu_matrix_intt_to(self.raw, dst_row, src.raw), "gpu_matrix_intt_to", self._L)
        self.is_ntt = False

    def add_scaled_block_from(self, src, scalar, dst_row, dst_col, src_row, src_col, rows, cols):
        """self[block] += src[block] * scalar (scalar 1x1; all EVAL) in one launch (extension gpu_matrix_add_scaled_block)"""
        check_status(self._L.gpu_matrix_add_scaled_block(self.raw, src.raw, scalar.raw, dst_row, dst_col, src_row, src_col, rows, cols),
                     "gpu_matrix_add_scaled_block", self._L)

    def mul_decompose_small(self, other, fused: bool | None = None):
        return self.mul_decompose(other, fused=fused, small=True)

    def mul_tensor_identity(self, other, identity_size):  # gpu_dcrt_poly.rs:1374-1390
        w = other.nrow
        assert self.ncol == w * identity_size
        parts = [self.slice(0, self.nrow, i * w, (i + 1) * w) * other for i in range(identity_size)]
        return parts[0].concat_columns(parts[1:])

    def mul_tensor_identity_decompose(self, other, identity_size):  # gpu_dcrt_poly.rs:1392-1412
        w = other.nrow * self.params.modulus_digits()
        assert self.ncol == w * identity_size
        parts = [self.slice(0, self.nrow, i * w, (i + 1) * w).mul_decompose(other) for i in range(identity_size)]
        return parts[0].concat_columns(parts[1:])

    # -- samplers -------------------------------------------------------
    @classmethod
    def sample_distribution(cls, params, nrow, ncol, dist: int, sigma: float, seed):  # gpu_dcrt_poly.rs:454-469
        out = cls.new_empty(params, nrow, ncol)
        if nrow == 0 or ncol == 0:
            return out
        check_status(out._L.gpu_matrix_sample_distribution(out.raw, int(dist), float(sigma), as_seed(seed)),
                     "gpu_matrix_sample_distribution", out._L)
        return out

    @classmethod
    def sample_distribution_columns(cls, params, nrow, total_ncol, col_start, col_len, dist: int, sigma: float, seed):
        """gpu_dcrt_poly.rs:471-507: columns [col_start, col_start+col_len) of the nrow x total_ncol matrix `seed` defines."""
        assert col_start + col_len <= total_ncol, "sample_distribution_columns range out of bounds"
        out = cls.new_empty(params, nrow, col_len)
        if nrow == 0 or col_len == 0:
            return out
        check_status(out._L.gpu_matrix_sample_distribution_columns(out.raw, int(dist), float(sigma), as_seed(seed),
                                                                   total_ncol, col_start),
                     "gpu_matrix_sample_distribution_columns", out._L)
        return out

    def gauss_samp_gq_arb_base(self, c: float, dgg_stddev: float, seed):  # gpu_dcrt_poly.rs:509-528 (consumes self)
        out = GpuDCRTPolyMatrix.new_empty(self.params, self.nrow * self.params.modulus_digits(), self.ncol)
        self.intt_all_in_place()
        check_status(self._L.gpu_matrix_gauss_samp_gq_arb_base(self.raw, self.params.base_bits, float(c), float(dgg_stddev),
                                                               as_seed(seed), out.raw),
                     "gpu_matrix_gauss_samp_gq_arb_base", self._L)
        return out

    @staticmethod
    def create_p1_covariance_cache(a_mat, b_mat, d_mat, sigma: float, s: float, dgg_stddev: float):  # :530-559
        return GpuP1CovarianceCache(a_mat, b_mat, d_mat, sigma, s, dgg_stddev)

    @classmethod
    def sample_p1_full_cached(cls, cache: "GpuP1CovarianceCache", tp2, seed):  # gpu_dcrt_poly.rs:561-574 (consumes tp2)
        out = cls.new_empty(tp2.params, tp2.nrow, tp2.ncol)
        if tp2.nrow == 0 or tp2.ncol == 0:
            return out
        tp2.intt_all_in_place()
        check_status(out._L.gpu_matrix_sample_p1_full_cached(cache.raw, tp2.raw, as_seed(seed), out.raw),
                     "gpu_matrix_sample_p1_full_cached", out._L)
        return out


# GpuMatrixSampleDist (gpu_dcrt_poly.rs:70-88)
DIST_UNIFORM, DIST_GAUSS, DIST_BIT, DIST_TERNARY = 0, 1, 2, 3


def as_seed(seed) -> GpuRngSeed:
    """GpuRngSeed::from_bytes (gpu.rs:51-61): 32 bytes -> four LE u64 words; also accepts 4 ints or one int (word 0)."""
    if isinstance(seed, GpuRngSeed):
        return seed
    if isinstance(seed, (bytes, bytearray)):
        assert len(seed) == 32
        words = [int.from_bytes(seed[8 * i:8 * i + 8], "little") for i in range(4)]
    elif isinstance(seed, int):
        words = [seed & (2**64 - 1), 0, 0, 0]
    else:
        words = [int(w) & (2**64 - 1) for w in seed]
        assert len(words) == 4
    out = GpuRngSeed()
    for i, w in enumerate(words):
        out.words[i] = w
    return out


class GpuP1CovarianceCache:
    """gpu_dcrt_poly.rs:56-68: owns the device tables of the perturbation covariance; freed on drop."""

    def __init__(self, a_mat, b_mat, d_mat, sigma, s, dgg_stddev):
        self._L = a_mat._L
        self._keep = (a_mat.params,)
        h = C.c_void_p()
        check_status(self._L.gpu_matrix_create_p1_covariance_cache(a_mat.raw, b_mat.raw, d_mat.raw, float(sigma), float(s),
                                                                   float(dgg_stddev), C.byref(h)),
                     "gpu_matrix_create_p1_covariance_cache", self._L)
        self.raw = h

    def destroy(self):
        if getattr(self, "raw", None):
            self._L.gpu_matrix_destroy_p1_covariance_cache(self.raw)
            self.raw = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass
