"""BGG+ public keys / encodings over the DCRT matrix path — the caller of the hot path for config C4
(SURVEY.md 8f-2).  Host mirror of src/bgg/public_key.rs and src/bgg/encoding.rs: same operator names,
same arithmetic, all of it through the gpu_* C ABI (nothing here touches residues).

    BggPublicKey * BggPublicKey   = A_x.mul_decompose(A_y)                              (public_key.rs:91-105)
    BggEncoding  * BggEncoding    = (c_x.mul_decompose(A_y) + c_y * x,  A_x.mul_decompose(A_y),  x*y)
                                                                                        (encoding.rs:125-144)

The reference computes G^-1(A_y) twice per encoding product (encoding.rs:131 and :140).  `fused=True`
stacks [c_x ; A_x] and issues ONE mul_decompose, so the digit NTTs — the dominant cost — are done once.

`ShardedBggEvaluator` runs a chain of products with the columns of every right operand sharded over the
ranks (mxx_b200/dist.py): each rank computes its columns of the new vector and public key, and one
all-gather per gate assembles them, because the next gate needs them whole as its LEFT operand.
"""
from __future__ import annotations

from .matrix import GpuDCRTPolyMatrix


class BggPublicKey:
    def __init__(self, matrix: GpuDCRTPolyMatrix, reveal_plaintext: bool = True):
        self.matrix = matrix
        self.reveal_plaintext = bool(reveal_plaintext)

    def __add__(self, other):  # public_key.rs:44-61
        return BggPublicKey(self.matrix + other.matrix, self.reveal_plaintext and other.reveal_plaintext)

    def __sub__(self, other):  # public_key.rs:63-82
        return BggPublicKey(self.matrix - other.matrix, self.reveal_plaintext and other.reveal_plaintext)

    def __mul__(self, other):  # public_key.rs:91-105
        return BggPublicKey(self.matrix.mul_decompose(other.matrix), self.reveal_plaintext and other.reveal_plaintext)

    def __eq__(self, other):
        return self.matrix == other.matrix and self.reveal_plaintext == other.reveal_plaintext


class BggEncoding:
    """vector: 1 x m (m = d * log_base_q), pubkey: d x m, plaintext: a 1x1 matrix (EVAL) or None"""

    def __init__(self, vector: GpuDCRTPolyMatrix, pubkey: BggPublicKey, plaintext: GpuDCRTPolyMatrix | None):
        self.vector, self.pubkey, self.plaintext = vector, pubkey, plaintext

    def __add__(self, other):  # encoding.rs:57-84
        pt = self.plaintext + other.plaintext if (self.plaintext is not None and other.plaintext is not None) else None
        return BggEncoding(self.vector + other.vector, self.pubkey + other.pubkey, pt)

    def __sub__(self, other):  # encoding.rs:86-113
        pt = self.plaintext - other.plaintext if (self.plaintext is not None and other.plaintext is not None) else None
        return BggEncoding(self.vector - other.vector, self.pubkey - other.pubkey, pt)

    def mul(self, other, fused: bool = True):  # encoding.rs:125-144
        if self.plaintext is None:
            raise ValueError("Unknown plaintext for the left-hand input of multiplication")
        a_y = other.pubkey.matrix
        if fused:
            stacked = self.vector.concat_rows([self.pubkey.matrix])          # [c_x ; A_x], (1 + d) x m
            prod = stacked.mul_decompose(a_y)                               # one G^-1(A_y)
            first = prod.slice_rows(0, self.vector.nrow)
            new_pk = prod.slice_rows(self.vector.nrow, prod.nrow)
        else:
            first = self.vector.mul_decompose(a_y)
            new_pk = self.pubkey.matrix.mul_decompose(a_y)
        new_vector = first + other.vector.mul_scalar(self.plaintext)
        pt = self.plaintext * other.plaintext if other.plaintext is not None else None
        reveal = self.pubkey.reveal_plaintext and other.pubkey.reveal_plaintext
        return BggEncoding(new_vector, BggPublicKey(new_pk, reveal), pt)

    __mul__ = mul


def encode(secret: GpuDCRTPolyMatrix, pubkey: BggPublicKey, plaintext: GpuDCRTPolyMatrix, gadget: GpuDCRTPolyMatrix,
           error: GpuDCRTPolyMatrix | None = None) -> BggEncoding:
    """c = s*A - x*(s*G) (+ e): the encoding relation the reference's tests check (src/bgg/mod.rs:98-111)."""
    vec = secret * pubkey.matrix - (secret * gadget).mul_scalar(plaintext)
    if error is not None:
        vec = vec + error
    return BggEncoding(vec, pubkey, plaintext)


class StackedEncoding:
    """A BGG+ encoding kept as ONE (1 + d) x m matrix [c ; A] (vector row(s) on top of the public key) plus its plaintext:
    the layout a product chain wants, because the stacked matrix IS the left operand of the next fused product — no
    concat before and no slices after each gate."""

    def __init__(self, stacked: GpuDCRTPolyMatrix, vec_rows: int, plaintext: GpuDCRTPolyMatrix | None):
        self.stacked, self.vec_rows, self.plaintext = stacked, int(vec_rows), plaintext

    @classmethod
    def from_encoding(cls, enc: BggEncoding) -> "StackedEncoding":
        return cls(enc.vector.concat_rows([enc.pubkey.matrix]), enc.vector.nrow, enc.plaintext)

    def to_encoding(self) -> BggEncoding:
        s = self.stacked
        return BggEncoding(s.slice_rows(0, self.vec_rows), BggPublicKey(s.slice_rows(self.vec_rows, s.nrow)), self.plaintext)


class ShardedBggEvaluator:
    """Chain of encoding products with right operands sharded by columns over the ranks."""

    def __init__(self, ops):
        self.ops = ops  # mxx_b200.dist.ColumnShardedOps

    def mul_gate(self, left: BggEncoding, right_vec_local: GpuDCRTPolyMatrix, right_pk_local: GpuDCRTPolyMatrix,
                 right_plaintext: GpuDCRTPolyMatrix | None, ncol: int) -> BggEncoding:
        """left: whole encoding (replicated); right_*_local: this rank's columns of c_y and A_y.
        Returns the whole product encoding on every rank (one all-gather of the stacked (1+d) x m block)."""
        return self.mul_gate_stacked(StackedEncoding.from_encoding(left), right_vec_local, right_pk_local, right_plaintext, ncol).to_encoding()

    def mul_gate_stacked(self, left: StackedEncoding, right_vec_local: GpuDCRTPolyMatrix, right_pk_local: GpuDCRTPolyMatrix,
                         right_plaintext: GpuDCRTPolyMatrix | None, ncol: int) -> StackedEncoding:
        """The same gate on stacked encodings: one fused mul_decompose on the local columns (no collective), the
        `+ c_y * x` tail applied in place to the vector rows (one launch), one in-place all-gather per row."""
        if left.plaintext is None:
            raise ValueError("Unknown plaintext for the left-hand input of multiplication")
        prod_local = self.ops.mul_decompose(left.stacked, right_pk_local)                   # no collective
        prod_local.add_scaled_block_from(right_vec_local, left.plaintext, 0, 0, 0, 0, left.vec_rows, prod_local.ncol)
        full = self.ops.all_gather_columns(prod_local, ncol)                               # the exchange step
        pt = left.plaintext * right_plaintext if right_plaintext is not None else None
        return StackedEncoding(full, left.vec_rows, pt)


class GridBggEvaluator:
    """The same chain on a 2-D (limb group x column block) partition (mxx_b200.dist.GridShardedOps).  Every matrix
    holds this rank's limbs only; the left encoding has all m columns, the right operands this rank's column block.
    Per gate: one all-gather of the coefficient-form A_y block along the limb axis, the local product, and — only when the
    columns are split too — one all-gather of the result rows along the column axis.

    The limb-axis exchange depends on the RIGHT operand only, and an evaluator knows its wires ahead of time: `prefetch`
    starts a gate's exchange (inverse transform + asynchronous in-place all-gather) so that it travels while the previous
    gate's product is computed; `mul_gate_stacked(..., gathered=handle)` then finds the towers already there."""

    def __init__(self, ops):
        self.ops = ops

    def prefetch(self, right_pk_local: GpuDCRTPolyMatrix):
        return self.ops.gather_towers_async(right_pk_local, side_stream=True)

    def mul_gate_stacked(self, left: StackedEncoding, right_vec_local: GpuDCRTPolyMatrix, right_pk_local: GpuDCRTPolyMatrix,
                         right_plaintext: GpuDCRTPolyMatrix | None, ncol: int, gathered=None) -> StackedEncoding:
        if left.plaintext is None:
            raise ValueError("Unknown plaintext for the left-hand input of multiplication")
        prod_local = self.ops.mul_decompose(left.stacked, right_pk_local, gathered=gathered)  # limb-axis all-gather
        prod_local.add_scaled_block_from(right_vec_local, left.plaintext, 0, 0, 0, 0, left.vec_rows, prod_local.ncol)
        full = self.ops.all_gather_columns(prod_local, ncol)                               # column-axis all-gather
        pt = left.plaintext * right_plaintext if right_plaintext is not None else None
        return StackedEncoding(full, left.vec_rows, pt)
