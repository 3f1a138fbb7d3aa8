"""CPU tests that pin the oracle (oracle/dcrt_oracle.c) before it is trusted.

Pins available in the reference tree: the literal CRT basis for
(n=4, depth 2, 17 bits), the eval-layout definition X[k] = a(psi^(2k+1)) with psi
from the least primitive root, and the reference's relational tests
(src/matrix/dcrt_poly.rs:515-799, src/matrix/gpu_dcrt_poly.rs:2077-2379).
"""
import numpy as np
import pytest

from oracle import oracle as orc

PARAM_SETS = [  # (n, depth, crt_bits, base_bits) used by the reference tests
    (4, 2, 17, 1),     # DCRTPolyParams::default, src/poly/dcrt/params.rs:73
    (4, 2, 17, 3),     # unaligned base, src/matrix/dcrt_poly.rs:683
    (4, 1, 52, 17),    # unaligned last digit, src/matrix/dcrt_poly.rs:704
    (128, 2, 17, 1),   # gpu_test_params, src/matrix/gpu_dcrt_poly.rs:1946
    (16, 4, 51, 17),
    (64, 3, 24, 12),   # bench moduli width, benches/bench_matrix_mul_cpu.rs:12
    (32, 2, 28, 14),   # production width, bench/*.csv
]


def plan_for(n, depth, bits):
    return orc.Plan(n, orc.prime_chain(n, bits, depth))


def test_prime_chain_literal_basis():
    # src/slot_transfer/bgg_poly_encoding_gpu.rs:1501 — GpuDCRTPolyParams::new(4, vec![131041, 131009], 1)
    assert orc.prime_chain(4, 17, 2) == [131041, 131009]


def test_prime_chain_properties():
    for n, bits, depth in [(16384, 24, 15), (16384, 51, 4), (65536, 28, 16), (4096, 50, 4)]:
        ch = orc.prime_chain(n, bits, depth)
        assert len(set(ch)) == depth and sorted(ch, reverse=True) == ch
        for p in ch:
            assert p < 2**bits and p % (2 * n) == 1 and orc.lib().orc_is_prime(p)
        # nothing skipped between consecutive members
        c = ch[0] + 2 * n
        while c < 2**bits:
            assert not orc.lib().orc_is_prime(c)
            c += 2 * n


def test_primitive_root_and_psi():
    sympy = pytest.importorskip("sympy")
    for p, n in [(131041, 4), (131009, 4), (orc.prime_chain(16384, 24, 1)[0], 16384), (orc.prime_chain(16, 51, 1)[0], 16)]:
        g = orc.lib().orc_least_primitive_root(p)
        assert g == sympy.primitive_root(p)
        psi = orc.lib().orc_psi(p, n)
        assert pow(psi, n, p) == p - 1 and psi == pow(g, (p - 1) // (2 * n), p)


def test_modmul_variants_agree():
    rng = np.random.default_rng(1)
    for bits in (17, 24, 28, 31, 51, 60, 62):
        q = orc.prime_chain(8, bits, 1)[0]
        for _ in range(200):
            a, b = int(rng.integers(0, q)), int(rng.integers(0, q))
            want = a * b % q
            assert orc.lib().orc_mulmod(a, b, q) == want
            assert orc.lib().orc_barrett_mulmod_check(a, b, q) == want
        for a, b in [(q - 1, q - 1), (0, q - 1), (1, 1)]:
            assert orc.lib().orc_barrett_mulmod_check(a, b, q) == a * b % q


@pytest.mark.parametrize("n,bits", [(4, 17), (8, 24), (64, 28), (256, 51), (128, 60)])
def test_ntt_three_ways_agree(n, bits):
    """definition == reference pipeline restatement == fast independent algorithm."""
    q = orc.prime_chain(n, bits, 1)[0]
    psi = orc.lib().orc_psi(q, n)
    plan = orc.Plan(n, [q])
    a = plan.uniform(1, 1, 7)
    naive = orc.ntt_naive(a[0, 0, 0], q, psi)
    ref = orc.ntt_ref(a[0, 0, 0], q, psi)
    fast = plan.ntt(a)[0, 0, 0]
    assert np.array_equal(naive, ref) and np.array_equal(naive, fast)
    # python big-int spot check of the definition
    k = n // 3
    x = pow(psi, 2 * k + 1, q)
    assert int(naive[k]) == sum(int(a[0, 0, 0, i]) * pow(x, i, q) for i in range(n)) % q
    assert np.array_equal(orc.intt_ref(ref, q, psi), a[0, 0, 0])
    assert np.array_equal(plan.intt(plan.ntt(a)), a)


@pytest.mark.parametrize("n,bits", [(16, 17), (128, 24), (512, 51)])
def test_eval_product_is_negacyclic_convolution(n, bits):
    q = orc.prime_chain(n, bits, 1)[0]
    plan = orc.Plan(n, [q])
    a, b = plan.uniform(1, 1, 1), plan.uniform(1, 1, 2)
    c = plan.intt(plan.mul(plan.ntt(a), plan.ntt(b)))
    assert np.array_equal(c[0, 0, 0], orc.negacyclic_schoolbook(a[0, 0, 0], b[0, 0, 0], q))


def test_ntt_large_spot_check():
    """n = 2^14: O(n) per-slot evaluation of the definition at a few slots."""
    n = 16384
    q = orc.prime_chain(n, 24, 1)[0]
    psi = orc.lib().orc_psi(q, n)
    plan = orc.Plan(n, [q])
    a = plan.uniform(1, 1, 3)
    X = plan.ntt(a)[0, 0, 0]
    coeffs = [int(v) for v in a[0, 0, 0]]
    for k in (0, 1, 5000, n - 1):
        x = pow(psi, 2 * k + 1, q)
        acc = 0
        for cval in reversed(coeffs):
            acc = (acc * x + cval) % q
        assert int(X[k]) == acc


def test_matrix_mul_matches_bigint():
    plan = plan_for(8, 3, 24)
    a, b = plan.uniform(2, 3, 11), plan.uniform(3, 4, 12)
    out = plan.mul(a, b)
    for l, q in enumerate(plan.moduli):
        A = a[:, :, l, :].astype(object)
        B = b[:, :, l, :].astype(object)
        want = np.einsum("ikn,kjn->ijn", A, B) % q
        assert np.array_equal(out[:, :, l, :].astype(object), want)


@pytest.mark.parametrize("n,depth,bits,bb", PARAM_SETS)
def test_gadget_times_decompose_is_identity(n, depth, bits, bb):
    """G * G^-1(M) == M (src/matrix/dcrt_poly.rs:515-600, 683-728;
    gpu_dcrt_poly.rs:2077-2163)."""
    plan = plan_for(n, depth, bits)
    size = 3
    ell = plan.ell(bb)
    M = plan.uniform(size, 2, 5)                    # EVAL form
    dec = plan.decompose(plan.intt(M), bb)          # COEFF digits
    assert dec.shape[0] == size * ell
    G = plan.ntt(plan.gadget(size, bb))
    assert np.array_equal(plan.mul(G, plan.ntt(dec)), M)
    # digit bound < 2^bb (gpu_dcrt_poly.rs:2251-2289)
    assert int(dec.max()) < (1 << bb)
    # every limb of a digit polynomial carries the same small integer
    for l in range(1, plan.L):
        assert np.array_equal(dec[:, :, l, :] % plan.moduli[l], dec[:, :, 0, :] % plan.moduli[l])


@pytest.mark.parametrize("n,depth,bits,bb", PARAM_SETS)
def test_decompose_digit_rule_python(n, depth, bits, bb):
    """Per-tower relation spelled out in gpu_dcrt_poly.rs:2494-2537."""
    plan = plan_for(n, depth, bits)
    dpt = plan.dpt(bb)
    src = plan.uniform(2, 2, 9)
    dec = plan.decompose(src, bb)
    ell = dpt * plan.L
    for row in range(2):
        for t, qt in enumerate(plan.moduli):
            acc = np.zeros((2, n), dtype=object)
            for dg in range(dpt):
                d = dec[row * ell + t * dpt + dg, :, 0, :].astype(object)
                width = max(0, min(bb, qt.bit_length() - dg * bb))
                assert (d < (1 << width)).all() if width else (d == 0).all()
                acc += d << (dg * bb)
            assert np.array_equal(acc, src[row, :, t, :].astype(object))


@pytest.mark.parametrize("n,depth,bits,bb", PARAM_SETS)
def test_small_gadget_and_mul_decompose_small(n, depth, bits, bb):
    """(A * G_small).mul_decompose_small(B) == A * B for B with small coefficients
    (src/matrix/dcrt_poly.rs:754-799, gpu_dcrt_poly.rs:2336-2379)."""
    plan = plan_for(n, depth, bits)
    d, c, r = 2, 3, 2
    dpt = plan.dpt(bb)
    A = plan.uniform(r, d, 21)
    rng = np.random.default_rng(3)
    small = rng.integers(0, min(plan.moduli), size=(d, c, 1, n), dtype=np.uint64)
    Bc = np.broadcast_to(small, (d, c, plan.L, n)).copy()   # same integer < q_0.. in every limb
    for l, q in enumerate(plan.moduli):
        Bc[:, :, l, :] %= q
    B = plan.ntt(Bc)
    Gs = plan.ntt(plan.gadget(d, bb, small=True))
    assert Gs.shape[1] == d * dpt
    lhs = plan.mul_decompose(plan.mul(A, Gs), B, bb, small=True)
    # B's integer coefficients are < min modulus, so limb 0 holds the integer itself
    assert np.array_equal(lhs, plan.mul(A, B))
    # G_small * small_decompose(B) == B (gpu_dcrt_poly.rs:2225-2247)
    dec = plan.decompose(Bc, bb, small=True)
    assert dec.shape[0] == d * dpt
    assert np.array_equal(plan.mul(Gs, plan.ntt(dec)), B)


@pytest.mark.parametrize("n,depth,bits,bb", PARAM_SETS)
def test_mul_decompose_equals_mul_of_decomposed(n, depth, bits, bb):
    """pk_a * pk_b == a.mul_decompose(b) (src/bgg/mod.rs:54-76) and
    chunked == full decomposition (dcrt_poly.rs:604-654)."""
    plan = plan_for(n, depth, bits)
    d, c, r = 2, 3, 2
    ell = plan.ell(bb)
    S = plan.uniform(r, d * ell, 31)
    other = plan.uniform(d, c, 32)
    full = plan.mul_decompose(S, other, bb)
    dec = plan.ntt(plan.decompose(plan.intt(other), bb))
    assert np.array_equal(full, plan.mul(S, dec))
    for j in range(c):  # column-chunked, width 1 (src/env.rs:148-154)
        col = plan.mul_decompose(S, other[:, j:j + 1], bb)
        assert np.array_equal(col, full[:, j:j + 1])
    # (A*G).mul_decompose(B) == A*B
    A = plan.uniform(r, d, 33)
    G = plan.ntt(plan.gadget(d, bb))
    assert np.array_equal(plan.mul_decompose(plan.mul(A, G), other, bb), plan.mul(A, other))


def test_ring_identities():
    """(p1-p2)+p2 == p1 etc. (src/poly/dcrt/poly.rs:631-720)."""
    plan = plan_for(16, 3, 28)
    a, b = plan.uniform(2, 2, 41), plan.uniform(2, 2, 42)
    assert np.array_equal(plan.add(plan.sub(a, b), b), a)
    s = plan.uniform(1, 1, 43)
    want = np.stack([np.stack([plan.mul(a[i:i + 1, j:j + 1], s)[0, 0] for j in range(2)]) for i in range(2)])
    assert np.array_equal(plan.mul_scalar(a, s), want)
