"""CPU checks of the CUDA block-level templates through the host emulation harness.

mxx_b200/libgpupoly_emu.so compiles the very templates the kernels are made of
(csrc/include/{ntt,arith}.cuh) as one-thread blocks and drives them with the launcher's
pass schedule.  Compared here, bit for bit, with the oracle.  This validates index math,
twiddle addressing, lazy-reduction bounds and the decompose/matmul geometry without a
GPU; the `-m gpu` tests then check the real kernels through the C ABI.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "mxx_b200", "libgpupoly_emu.so")
u64p = C.POINTER(C.c_uint64)


@pytest.fixture(scope="module")
def emu():
    if not os.path.exists(EMU):
        import __graft_entry__ as ge

        ge.build()
    lib = C.CDLL(EMU)
    lib.emu_ntt.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, u64p, C.c_size_t, C.c_int, C.c_int]
    lib.emu_decompose.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, u64p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                  C.c_uint32, C.c_int, C.c_int, u64p, C.c_int]
    lib.emu_set_f64.argtypes = [C.c_int]
    lib.emu_set_f64.restype = None
    lib.emu_set_fuse_top.argtypes = [C.c_int]
    lib.emu_set_fuse_top.restype = None
    lib.emu_f64_blocks.restype = C.c_size_t
    lib.emu_set_cluster.argtypes = [C.c_int]
    lib.emu_set_cluster.restype = None
    lib.emu_matmul.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, u64p, u64p, u64p] + [C.c_uint32] * 8 + [C.c_int]
    return lib


def p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def bitrev_perm(n):
    bits = n.bit_length() - 1
    idx = np.arange(n)
    out = np.zeros(n, dtype=np.int64)
    for b in range(bits):
        out |= ((idx >> b) & 1) << (bits - 1 - b)
    return out


def to_dev(eval_nat):
    """natural-order EVAL (boundary layout) -> device order (slot s holds X[bitrev(s)])"""
    n = eval_nat.shape[-1]
    return np.ascontiguousarray(eval_nat[..., bitrev_perm(n)])


def plan(n, depth, bits):
    return orc.Plan(n, orc.prime_chain(n, bits, depth))


CASES = [
    # (n, depth, bits, logb_cap, force_wide)
    (4, 2, 17, 0, 0),
    (8, 1, 20, 0, 0),
    (64, 3, 24, 0, 0),
    (1024, 2, 28, 0, 0),
    (4096, 2, 29, 0, 0),
    (2048, 2, 51, 0, 0),
    (256, 2, 61, 0, 0),
    (1024, 2, 24, 0, 1),
    # blocks smaller than the transform: exercises the global top passes
    (1024, 2, 28, 6, 0),
    (1024, 1, 28, 5, 0),
    (2048, 1, 28, 3, 0),
    (4096, 2, 51, 7, 0),
    (512, 1, 51, 8, 0),
    (256, 1, 24, 1, 0),
    # compile-time-planned kernels (ntt_fast.cuh): NOCSUB (24/26-bit, 51-bit) and csub (28/29-bit, 61-bit)
    (4096, 2, 24, 0, 0),
    (4096, 1, 29, 0, 0),
    (8192, 1, 26, 0, 0),
    (8192, 1, 28, 0, 0),
    (16384, 2, 24, 0, 0),
    (16384, 1, 29, 0, 0),
    (32768, 1, 24, 0, 0),
    (32768, 1, 28, 0, 0),
    (65536, 1, 24, 0, 0),   # 2^14 blocks under a radix-4 global pass
    (65536, 1, 29, 0, 0),
    (4096, 1, 51, 0, 0),
    (8192, 1, 61, 0, 0),
    (16384, 1, 51, 0, 0),
    (16384, 1, 61, 0, 0),
    (32768, 1, 51, 0, 0),   # 2^13 blocks under a radix-4 global pass
    (65536, 1, 61, 0, 0),
    (16384, 1, 55, 0, 0),   # u64 integer NOCSUB plan above the FP64 range
    (32768, 1, 55, 0, 0),
    (16384, 1, 24, 0, 1),   # u64 storage forced for a small modulus
    (16384, 1, 24, 14, 0),  # same ring through the generic kernel
]


@pytest.fixture
def emu_f64(emu):
    """drive the FP64-pipe kernels (modarith.cuh F64) the library picks for u64 contexts with moduli < 2^51"""
    emu.emu_set_f64(1)
    before = emu.emu_f64_blocks()
    yield emu
    emu.emu_set_f64(0)
    assert emu.emu_f64_blocks() > before, "the F64 kernels were not exercised"


F64_CASES = [
    # (n, depth, bits, force_wide)
    (4096, 2, 51, 0),
    (8192, 1, 51, 0),
    (8192, 2, 40, 0),
    (16384, 2, 51, 0),
    (16384, 1, 31, 0),
    (32768, 1, 51, 0),    # 2^13 blocks under an integer radix-4 global pass
    (65536, 1, 51, 0),
    (16384, 1, 24, 1),    # u64 storage forced for a small modulus
]


@pytest.mark.parametrize("n,depth,bits,fw", F64_CASES)
def test_emu_f64_ntt_matches_oracle(emu_f64, n, depth, bits, fw):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    assert int(mod.max()) < 1 << 51
    a = pl.uniform(1, 3, seed=23 + n)
    want = to_dev(pl.ntt(a))
    d = a.copy()
    assert emu_f64.emu_ntt(fw, n.bit_length() - 1, depth, p(mod), p(d), 3, 0, 0) == 0
    assert np.array_equal(d, want)
    assert emu_f64.emu_ntt(fw, n.bit_length() - 1, depth, p(mod), p(d), 3, 1, 0) == 0
    assert np.array_equal(d, a)


def test_emu_f64_mixed_modulus_sizes(emu_f64):
    """a wide context mixing 20-, 35- and 51-bit limbs: per-limb 1/q, fold count from the largest modulus"""
    n = 4096
    moduli = orc.prime_chain(n, 20, 1) + orc.prime_chain(n, 51, 1) + orc.prime_chain(n, 35, 1)
    pl = orc.Plan(n, moduli)
    mod = np.array(moduli, dtype=np.uint64)
    a = pl.uniform(1, 2, seed=77)
    a[0, 1] = (mod - 1)[:, None]
    d = a.copy()
    assert emu_f64.emu_ntt(0, 12, 3, p(mod), p(d), 2, 0, 0) == 0
    assert np.array_equal(d, to_dev(pl.ntt(a)))
    assert emu_f64.emu_ntt(0, 12, 3, p(mod), p(d), 2, 1, 0) == 0
    assert np.array_equal(d, a)
    pl16 = orc.Plan(16, [orc.prime_chain(16, b, 1)[0] for b in (20, 51, 35)])
    mod16 = np.array(pl16.moduli, dtype=np.uint64)
    A, B = pl16.uniform(2, 9, seed=1), pl16.uniform(9, 4, seed=2)
    A[1] = (mod16 - 1)[:, None]
    want = pl16.mul(A, B)
    out = np.zeros_like(want)
    assert emu_f64.emu_matmul(0, 4, 3, p(mod16), p(out), p(A), p(B), 2, 9, 4, 4, 0, 4, 0, 0, 1) == 0
    assert np.array_equal(out, want)


def test_emu_f64_ntt_extreme_values(emu_f64):
    """all coefficients q-1 / 0 / alternating: the largest un-renormalised legs the FP64 butterflies see"""
    for n, depth, bits in [(4096, 2, 51), (16384, 1, 51), (32768, 1, 51), (8192, 1, 33)]:
        pl = plan(n, depth, bits)
        mod = np.array(pl.moduli, dtype=np.uint64)
        a = np.empty((1, 3, depth, n), dtype=np.uint64)
        a[0, 0] = (mod - 1)[:, None]
        a[0, 1] = 0
        a[0, 2] = (mod - 1)[:, None]
        a[0, 2, :, ::2] = 0
        want = to_dev(pl.ntt(a))
        d = a.copy()
        assert emu_f64.emu_ntt(0, n.bit_length() - 1, depth, p(mod), p(d), 3, 0, 0) == 0
        assert np.array_equal(d, want)
        e = to_dev(a)
        want_c = pl.intt(a)
        assert emu_f64.emu_ntt(0, n.bit_length() - 1, depth, p(mod), p(e), 3, 1, 0) == 0
        assert np.array_equal(e, want_c)


@pytest.mark.parametrize("n,depth,bits,bb,small", [(4096, 1, 51, 17, 0), (8192, 2, 51, 13, 0), (4096, 2, 40, 20, 1), (32768, 1, 51, 17, 0)])
def test_emu_f64_decompose_matches_oracle(emu_f64, n, depth, bits, bb, small):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    rows, cols = 2, 3
    src = pl.uniform(rows, cols, seed=5)
    full = to_dev(pl.ntt(pl.decompose(src, bb, bool(small))))
    ell = pl.ell(bb, bool(small))
    for c0, w in [(0, cols), (1, 2)]:
        out = np.zeros((rows * ell, w, depth, n), dtype=np.uint64)
        assert emu_f64.emu_decompose(0, n.bit_length() - 1, depth, p(mod), p(src), rows, cols, c0, w, bb, small, 1, p(out), 0) == 0
        assert np.array_equal(out, np.ascontiguousarray(full[:, c0:c0 + w]))


@pytest.mark.parametrize("n,depth,bits,cap,fw", CASES)
def test_emu_ntt_matches_oracle(emu, n, depth, bits, cap, fw):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    a = pl.uniform(1, 3, seed=11 + n)
    want = to_dev(pl.ntt(a))
    d = a.copy()
    assert emu.emu_ntt(fw, n.bit_length() - 1, depth, p(mod), p(d), 3, 0, cap) == 0
    assert np.array_equal(d, want)
    # inverse of the device-order EVAL data gives back the coefficients
    assert emu.emu_ntt(fw, n.bit_length() - 1, depth, p(mod), p(d), 3, 1, cap) == 0
    assert np.array_equal(d, a)


def test_emu_ntt_extreme_values(emu):
    """all coefficients q-1 (largest lazy values) and all zero"""
    for n, depth, bits in [(4096, 2, 29), (1024, 2, 61), (1024, 2, 51), (16384, 1, 26), (16384, 1, 29), (16384, 1, 57), (16384, 1, 61)]:
        pl = plan(n, depth, bits)
        mod = np.array(pl.moduli, dtype=np.uint64)
        a = np.empty((1, 2, depth, n), dtype=np.uint64)
        a[0, 0] = (mod - 1)[:, None]
        a[0, 1] = 0
        want = to_dev(pl.ntt(a))
        d = a.copy()
        assert emu.emu_ntt(0, n.bit_length() - 1, depth, p(mod), p(d), 2, 0, 0) == 0
        assert np.array_equal(d, want)
        e = to_dev(a)  # treat the same values as EVAL data
        want_c = pl.intt(a)
        assert emu.emu_ntt(0, n.bit_length() - 1, depth, p(mod), p(e), 2, 1, 0) == 0
        assert np.array_equal(e, want_c)


DEC_CASES = [
    # (n, depth, bits, bb, small, cap)
    (4, 2, 17, 1, 0, 0),
    (4, 2, 17, 3, 0, 0),
    (4, 1, 52, 17, 0, 0),  # unaligned last digit (src/matrix/dcrt_poly.rs:683-728)
    (128, 2, 17, 1, 0, 0),
    (256, 3, 24, 12, 0, 0),
    (256, 2, 51, 17, 0, 0),
    (256, 3, 24, 12, 1, 0),
    (512, 2, 28, 14, 0, 5),
    (64, 2, 51, 20, 1, 0),
    (4096, 2, 24, 12, 0, 0),   # fast fused digit + NTT kernels
    (8192, 1, 28, 14, 0, 0),
    (4096, 1, 51, 17, 0, 0),
    (4096, 2, 24, 12, 1, 0),
    # rings above one block: digits extracted inside the first global pass (radix 2 / 4 / 8 / 16)
    (128, 2, 24, 8, 1, 6),
    (256, 1, 51, 17, 0, 6),
    (1024, 2, 24, 12, 0, 7),
    (2048, 1, 28, 14, 0, 5),
    (32768, 1, 51, 17, 0, 0),
    (65536, 1, 28, 14, 1, 0),
]


@pytest.mark.parametrize("n,depth,bits,bb,small,cap", DEC_CASES)
@pytest.mark.parametrize("to_eval", [0, 1])
def test_emu_decompose_matches_oracle(emu, n, depth, bits, bb, small, cap, to_eval):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    rows, cols = 2, 3
    src = pl.uniform(rows, cols, seed=5)
    full = pl.decompose(src, bb, bool(small))
    if to_eval:
        full = to_dev(pl.ntt(full))
    ell = pl.ell(bb, bool(small))
    # whole matrix, then a column chunk [1,3)
    for c0, w in [(0, cols), (1, 2)]:
        out = np.zeros((rows * ell, w, depth, n), dtype=np.uint64)
        st = emu.emu_decompose(0, n.bit_length() - 1, depth, p(mod), p(src), rows, cols, c0, w, bb, small, to_eval, p(out), cap)
        assert st == 0
        assert np.array_equal(out, np.ascontiguousarray(full[:, c0:c0 + w]))


def test_emu_decompose_unfused_top_pass(emu):
    """MXX_DECOMP_FUSE_TOP=0 path: digit kernel, then every global pass"""
    emu.emu_set_fuse_top(0)
    try:
        test_emu_decompose_matches_oracle(emu, 512, 2, 28, 14, 0, 5, 1)
        test_emu_decompose_matches_oracle(emu, 256, 1, 51, 17, 0, 6, 1)
    finally:
        emu.emu_set_fuse_top(1)


MM_CASES = [
    # (n, depth, bits, r, k, c, variant, fold)
    (16, 2, 24, 1, 5, 7, 0, 0),
    (16, 2, 29, 1, 30, 6, 0, 0),
    (16, 2, 29, 3, 9, 5, 1, 0),
    (16, 2, 29, 5, 9, 6, 2, 0),
    (16, 2, 29, 2, 40, 3, 0, 3),   # forced early folds
    (16, 2, 51, 1, 5, 7, 0, 0),
    (16, 2, 61, 3, 21, 5, 1, 0),   # 61-bit: 16 products per 128-bit accumulator
    (16, 2, 61, 5, 9, 6, 2, 2),
    (2, 1, 20, 2, 3, 2, 3, 0),     # ring smaller than a vector: scalar variant
    (4, 2, 17, 2, 3, 2, 3, 0),
]


@pytest.mark.parametrize("n,depth,bits,r,k,c,variant,fold", MM_CASES)
def test_emu_matmul_matches_oracle(emu, n, depth, bits, r, k, c, variant, fold):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    A = pl.uniform(r, k, seed=1)
    B = pl.uniform(k, c, seed=2)
    want = pl.mul(A, B)
    out = np.zeros_like(want)
    assert emu.emu_matmul(0, n.bit_length() - 1, depth, p(mod), p(out), p(A), p(B), r, k, c, c, 0, c, 0, fold, variant) == 0
    assert np.array_equal(out, want)
    # a column window of B into a column window of a wider output
    if c >= 3:
        wide = np.full((r, c + 2, depth, n), 7, dtype=np.uint64)
        assert emu.emu_matmul(0, n.bit_length() - 1, depth, p(mod), p(wide), p(A), p(B), r, k, c - 1, c, 1, c + 2, 2, fold, variant) == 0
        assert np.array_equal(wide[:, 2:c + 1], want[:, 1:])
        assert np.all(wide[:, :2] == 7) and np.all(wide[:, c + 1:] == 7)


F64_MM_CASES = [
    # (n, depth, bits, r, k, c, variant, fold)
    (16, 2, 51, 1, 30, 6, 0, 0),
    (16, 2, 51, 1, 5, 7, 0, 0),
    (16, 2, 51, 3, 21, 5, 1, 0),
    (16, 2, 51, 5, 9, 6, 2, 0),
    (16, 2, 40, 2, 40, 3, 1, 0),    # smaller moduli: more terms between renormalisations
    (16, 1, 31, 1, 70, 4, 3, 0),
    (16, 2, 51, 2, 24, 4, 1, 1),    # renormalise after every term
]


@pytest.mark.parametrize("n,depth,bits,r,k,c,variant,fold", F64_MM_CASES)
def test_emu_f64_matmul_matches_oracle(emu_f64, n, depth, bits, r, k, c, variant, fold):
    test_emu_matmul_matches_oracle(emu_f64, n, depth, bits, r, k, c, variant, fold)


def test_emu_f64_matmul_worst_case_accumulation(emu_f64):
    """every operand q-1 / mixed 0, 1, q-1: the largest terms and quotient-estimate errors of the FP64 product"""
    for bits, n in [(51, 16), (50, 16), (33, 16)]:
        pl = plan(n, 1, bits)
        q = pl.moduli[0]
        mod = np.array(pl.moduli, dtype=np.uint64)
        k = 70
        A = np.full((2, k, 1, n), q - 1, dtype=np.uint64)
        B = np.full((k, 4, 1, n), q - 1, dtype=np.uint64)
        rng = np.random.default_rng(bits)
        A[1] = rng.choice(np.array([0, 1, q - 1, q // 2, q // 2 + 1], dtype=np.uint64), size=(k, 1, n))
        B[:, 1] = rng.choice(np.array([0, 1, q - 1, q // 2, q // 2 + 1], dtype=np.uint64), size=(k, 1, n))
        want = pl.mul(A, B)
        out = np.zeros_like(want)
        assert emu_f64.emu_matmul(0, n.bit_length() - 1, 1, p(mod), p(out), p(A), p(B), 2, k, 4, 4, 0, 4, 0, 0, 1) == 0
        assert np.array_equal(out, want)


def test_emu_matmul_worst_case_accumulation(emu):
    """every operand q-1: the largest lazily accumulated sums"""
    for bits, n in [(29, 16), (61, 16), (24, 16)]:
        pl = plan(n, 1, bits)
        mod = np.array(pl.moduli, dtype=np.uint64)
        k = 70
        A = np.full((1, k, 1, n), pl.moduli[0] - 1, dtype=np.uint64)
        B = np.full((k, 4, 1, n), pl.moduli[0] - 1, dtype=np.uint64)
        want = pl.mul(A, B)
        out = np.zeros_like(want)
        assert emu.emu_matmul(0, n.bit_length() - 1, 1, p(mod), p(out), p(A), p(B), 1, k, 4, 4, 0, 4, 0, 0, 0) == 0
        assert np.array_equal(out, want)


# ---- cluster / DSMEM kernels (ntt_cluster.cuh): rings of 2, 4, 8 blocks ------------------------------------------------
# With the cluster path on (the launcher's default) the ring goes through push -> local (forward) and local -> pull
# (inverse) phases; with it off, through the round-1 global passes + block kernel.  Both must equal the oracle.
CLUSTER_CASES = [
    # (n, depth, bits, f64)
    (1 << 15, 2, 28, 0),   # u32, 2 blocks of 2^14
    (1 << 15, 1, 24, 0),   # NOCSUB
    (1 << 16, 1, 29, 0),   # 4 blocks
    (1 << 17, 1, 26, 0),   # 8 blocks, NOCSUB
    (1 << 14, 2, 61, 0),   # u64 integer, 2 blocks of 2^13
    (1 << 15, 1, 55, 0),   # u64 integer NOCSUB, 4 blocks
    (1 << 16, 1, 59, 0),   # 8 blocks
    (1 << 14, 2, 51, 1),   # FP64-pipe butterflies
    (1 << 15, 1, 51, 1),
    (1 << 16, 1, 45, 1),
]


@pytest.mark.parametrize("n,depth,bits,f64", CLUSTER_CASES)
def test_emu_cluster_ntt_matches_oracle_and_global_pass_path(emu, n, depth, bits, f64):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    logn = n.bit_length() - 1
    a = pl.uniform(1, 2, seed=31 + n)
    a[0, 1, :, : n // 2] = (mod - 1)[:, None]  # half of one polynomial at the largest residue
    want = to_dev(pl.ntt(a))
    emu.emu_set_f64(f64)
    try:
        before = emu.emu_f64_blocks()
        for cluster in (1, 0):
            emu.emu_set_cluster(cluster)
            d = a.copy()
            assert emu.emu_ntt(0, logn, depth, p(mod), p(d), 2, 0, 0) == 0
            assert np.array_equal(d, want), f"forward, cluster={cluster}"
            assert emu.emu_ntt(0, logn, depth, p(mod), p(d), 2, 1, 0) == 0
            assert np.array_equal(d, a), f"inverse, cluster={cluster}"
        if f64:
            assert emu.emu_f64_blocks() > before
    finally:
        emu.emu_set_cluster(1)
        emu.emu_set_f64(0)


@pytest.mark.parametrize("n,depth,bits,bb,small,f64", [(1 << 15, 2, 28, 14, 0, 0), (1 << 15, 1, 24, 12, 1, 0), (1 << 16, 1, 28, 10, 0, 0),
                                                     (1 << 14, 2, 51, 17, 0, 1), (1 << 15, 1, 51, 17, 0, 1), (1 << 14, 1, 61, 21, 0, 0)])
def test_emu_cluster_decompose_matches_oracle(emu, n, depth, bits, bb, small, f64):
    pl = plan(n, depth, bits)
    mod = np.array(pl.moduli, dtype=np.uint64)
    rows, cols = 1, 3
    src = pl.uniform(rows, cols, seed=9)
    full = to_dev(pl.ntt(pl.decompose(src, bb, bool(small))))
    ell = pl.ell(bb, bool(small))
    emu.emu_set_f64(f64)
    try:
        for cluster in (1, 0):
            emu.emu_set_cluster(cluster)
            out = np.zeros((rows * ell, 2, depth, n), dtype=np.uint64)
            assert emu.emu_decompose(0, n.bit_length() - 1, depth, p(mod), p(src), rows, cols, 1, 2, bb, small, 1, p(out), 0) == 0
            assert np.array_equal(out, np.ascontiguousarray(full[:, 1:3])), f"cluster={cluster}"
    finally:
        emu.emu_set_cluster(1)
        emu.emu_set_f64(0)
