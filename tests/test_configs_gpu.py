"""BASELINE.json configs other than the bench line (C1-J, C3-b, C3-c, C4, C5 of SURVEY.md 8) as GPU parity
tests: the CUDA path through the C ABI against the CPU oracle where the oracle finishes in seconds, and
through exact size-independent relations at the full sizes (the relations are the reference's own:
src/bgg/mod.rs:54-111, src/sampler/trapdoor/gpu.rs tests "public_matrix * preimage == target")."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

_cache = {}


def setup(n, depth, bits, bb):
    import mxx_b200 as mx

    key = (n, depth, bits, bb)
    if key not in _cache:
        moduli = mx.standard_moduli(n, depth, bits)
        _cache[key] = (mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli))
    return _cache[key]


def G(params, host, is_ntt=True):
    import mxx_b200 as mx

    return mx.GpuDCRTPolyMatrix.from_rns(params, host, is_ntt)


# ---- C1-J: small square products, ring 2^12, four ~50-bit limbs ------------------------------------
@pytest.mark.parametrize("size", [8, 16])
def test_c1j_square_products_match_oracle(size):
    params, plan = setup(4096, 4, 50, 10)
    A, B = plan.uniform(size, size, seed=201), plan.uniform(size, size, seed=202)
    got = (G(params, A) * G(params, B)).to_rns()
    assert np.array_equal(got, plan.mul(A, B))


# ---- C3-b / C3-c: mul_decompose with d > 1 ------------------------------------------------------------
@pytest.mark.parametrize("n,depth,bits,bb,r,d,c", [(16384, 15, 24, 12, 4, 4, 6), (16384, 4, 51, 17, 2, 2, 24), (16384, 15, 24, 12, 1, 4, 5)])
def test_c3_mul_decompose_rows_and_blocks(n, depth, bits, bb, r, d, c):
    params, plan = setup(n, depth, bits, bb)
    ell = plan.ell(bb)
    S, other = plan.uniform(r, d * ell, seed=211), plan.uniform(d, c, seed=212)
    gS, gO = G(params, S), G(params, other)
    out = gS.mul_decompose(gO)
    # sampled columns against the oracle (the full product is minutes of CPU time at d = 4)
    for j in (0, c - 1):
        want = plan.mul_decompose(S, other[:, j:j + 1], bb)
        assert np.array_equal(out.slice_columns(j, j + 1).to_rns(), want)
    # exact relations on the whole result
    import mxx_b200 as mx

    assert mx.GpuDCRTPolyMatrix.gadget_matrix(params, d).mul_decompose(gO) == gO
    assert out == gS.mul_decompose(gO, fused=False)


# ---- C4: BGG+ encoding products --------------------------------------------------------------------------
def _bgg_setup(params, plan, d, seed):
    import mxx_b200 as mx
    from mxx_b200.bgg import BggPublicKey, encode

    ell = params.modulus_digits()
    m = d * ell
    gadget = mx.GpuDCRTPolyMatrix.gadget_matrix(params, d)
    secret = G(params, plan.uniform(1, d, seed=seed))
    encs = []
    for i in range(3):
        pk = BggPublicKey(G(params, plan.uniform(d, m, seed=seed + 10 + i)))
        x = G(params, plan.uniform(1, 1, seed=seed + 20 + i))
        encs.append(encode(secret, pk, x, gadget))
    return secret, gadget, encs


def _relation_holds(secret, gadget, enc):
    """c == s*A - x*(s*G): the (error-free) encoding relation, exact in Z_q[x]/(x^n+1)"""
    return enc.vector == secret * enc.pubkey.matrix - (secret * gadget).mul_scalar(enc.plaintext)


@pytest.mark.parametrize("n,depth,bits,bb,d", [(256, 2, 24, 12, 1), (256, 2, 24, 12, 2), (1024, 2, 51, 17, 1)])
def test_bgg_mul_matches_oracle_and_relation(n, depth, bits, bb, d):
    params, plan = setup(n, depth, bits, bb)
    secret, gadget, (ex, ey, ez) = _bgg_setup(params, plan, d, seed=300)
    fused, plain = ex.mul(ey, fused=True), ex.mul(ey, fused=False)
    assert fused.vector == plain.vector and fused.pubkey.matrix == plain.pubkey.matrix
    # the public key product is S.mul_decompose(other) on the CPU oracle (src/bgg/mod.rs:54-76)
    want_pk = plan.mul_decompose(ex.pubkey.matrix.to_rns(), ey.pubkey.matrix.to_rns(), bb)
    assert np.array_equal(fused.pubkey.matrix.to_rns(), want_pk)
    want_vec = plan.mul_decompose(ex.vector.to_rns(), ey.pubkey.matrix.to_rns(), bb)
    first = G(params, want_vec) + ey.vector.mul_scalar(ex.plaintext)
    assert fused.vector == first
    assert _relation_holds(secret, gadget, fused)
    assert _relation_holds(secret, gadget, fused.mul(ez))  # depth 2
    assert _relation_holds(secret, gadget, (ex + ey).mul(ez - ex))


@pytest.mark.parametrize("depth,bits,bb", [(8, 51, 17), (16, 28, 14)])
def test_c4_full_size_bgg_chain_relation(depth, bits, bb):
    """n = 2^15, d = 1: a chain of encoding products keeps the encoding relation exactly"""
    params, plan = setup(1 << 15, depth, bits, bb)
    secret, gadget, (ex, ey, ez) = _bgg_setup(params, plan, 1, seed=400)
    acc = ex
    for other in (ey, ez, ey):
        acc = acc.mul(other)
        assert _relation_holds(secret, gadget, acc)
    # one output polynomial of the first gate against the oracle
    first = ex.mul(ey)
    j = 3
    want = plan.mul_decompose(ex.pubkey.matrix.to_rns(), ey.pubkey.matrix.to_rns()[:, j:j + 1], bb)
    assert np.array_equal(first.pubkey.matrix.slice_columns(j, j + 1).to_rns(), want)


# ---- C5: preimage-shaped solve A*x = u, verified exactly --------------------------------------------------
def test_c5_preimage_shaped_solution_is_exact():
    """bench_preimage_gpu shapes (benches/bench_preimage_gpu.rs:34-53): n = 2^14, 10 x 24-bit limbs, base 2^12,
    d = 1, 50 target columns.  With A = [a | 1 | G - (a r + e)] (src/sampler/trapdoor/gpu.rs:202-215) and a
    perturbation p, x = p + [r; e; 1] * z with z = G^-1(u - A p) satisfies A x = u exactly — the assembly the
    sampler performs (trapdoor/gpu.rs:276-368) with the deterministic z the reference tests also accept
    (z_plain = target.decompose(), trapdoor/gpu.rs:652)."""
    import mxx_b200 as mx

    n, depth, bits, bb, d, cols = 1 << 14, 10, 24, 12, 1, 50
    params, plan = setup(n, depth, bits, bb)
    ell = params.modulus_digits()
    a = G(params, plan.uniform(d, d, seed=501))
    r = G(params, plan.uniform(d, d * ell, seed=502))
    e = G(params, plan.uniform(d, d * ell, seed=503))
    gadget = mx.GpuDCRTPolyMatrix.gadget_matrix(params, d)
    one = mx.GpuDCRTPolyMatrix.identity(params, d)
    A = a.concat_columns([one, gadget - (a * r + e)])          # d x (ell + 2) d
    assert A.size() == (d, (ell + 2) * d)
    u = G(params, plan.uniform(d, cols, seed=504))
    p = G(params, plan.uniform((ell + 2) * d, cols, seed=505))  # perturbation stand-in
    z = (u - A * p).decompose()                                 # (d ell) x cols, G z = u - A p
    x = p + (r * z).concat_rows([e * z, z])
    assert x.size() == ((ell + 2) * d, cols)
    assert A * x == u
    # and the decomposition that carries it is the oracle's
    src = (u - A * p).slice_columns(0, 1).to_rns()
    want = plan.ntt(plan.decompose(plan.intt(src), bb))
    assert np.array_equal(z.slice_columns(0, 1).to_rns(), want)
