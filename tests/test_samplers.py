"""Samplers of the trapdoor path: random streams, the four matrix
distributions, the G-lattice sampler and the perturbation p1.

CPU (`-m "not gpu"`): the oracle against published known answers (RFC 8439 §2.3.2 ChaCha20
block, draft-irtf-cfrg-xchacha §2.2.1 HChaCha20) and against golden vectors produced by the
reference's own CUDA library; the product's per-element code (host emulation harness)
against the oracle; the relations the reference's tests assert
(src/matrix/gpu_dcrt_poly.rs:2383-2540, src/sampler/gpu.rs tests).
GPU (`-m gpu`): libgpupoly.so through the C ABI against the oracle, the golden vectors and
the same relations, and the full trapdoor -> preimage flow A x = u.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "mxx_b200", "libgpupoly_emu.so")
GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_cuda_sampler_vectors.npz")
u64p = C.POINTER(C.c_uint64)
SIGMA = 4.578  # the reference tests' sigma (src/sampler/trapdoor/gpu.rs:496)

DIST_UNIFORM, DIST_GAUSS, DIST_BIT, DIST_TERNARY = 0, 1, 2, 3
CASES = [(128, 2, 16, 8), (64, 3, 24, 12), (64, 2, 51, 17), (32, 2, 28, 1)]  # (n, depth, bits, base_bits)


def p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


def plan_for(n, depth, bits):
    return orc.Plan(n, orc.prime_chain(n, bits, depth))


def centred(m, moduli):
    """(rows, cols, L, n) residues -> signed representatives per limb"""
    q = np.array(moduli, dtype=np.int64).reshape(1, 1, -1, 1)
    v = m.astype(np.int64)
    return np.where(v > q // 2, v - q, v)


def gadget_recompose(plan, z_coeff, bb, rows):
    """sum_dg b^dg * z[row*ell + t*dpt + dg] mod q_t, per tower t — the relation of gpu_dcrt_poly.rs:2494-2537."""
    dpt, L = plan.dpt(bb), plan.L
    ell = dpt * L
    cols = z_coeff.shape[1]
    out = np.zeros((rows, cols, L, plan.n), dtype=object)
    for r in range(rows):
        for t in range(L):
            q = plan.moduli[t]
            acc = np.zeros((cols, plan.n), dtype=object)
            for dg in range(dpt):
                acc = (acc + pow(2, bb * dg, q) * z_coeff[r * ell + t * dpt + dg, :, t, :].astype(object)) % q
            out[r, :, t, :] = acc
    return out


# ---------------------------------------------------------------- oracle: known answers
def test_chacha20_block_rfc8439_known_answer():
    key = [int.from_bytes(bytes(range(4 * i, 4 * i + 4)), "little") for i in range(8)]
    got = orc.chacha20_block(key, [1, 0x09000000, 0x4A000000, 0])
    want = [0xE4E7F110, 0x15593BD1, 0x1FDD0F50, 0xC47120A3, 0xC7F4D1C7, 0x0368C033, 0x9AAA2204, 0x4E6CD4C3,
            0x466482D2, 0x09AA9F07, 0x05D7C214, 0xA2028BD9, 0xD19C12B5, 0xB94E16DE, 0xE883D0CB, 0x4E3C50A2]
    assert [int(x) for x in got] == want


def test_hchacha20_draft_known_answer():
    key = [int.from_bytes(bytes(range(4 * i, 4 * i + 4)), "little") for i in range(8)]
    nonce = bytes.fromhex("000000090000004a0000000031415927")
    got = orc.hchacha20(key, [int.from_bytes(nonce[4 * i:4 * i + 4], "little") for i in range(4)])
    want = bytes.fromhex("82413b4227b27bfed30e42508a877d73a0f9e4d58a74a853c12ec41326d3ecdc")
    assert b"".join(int(x).to_bytes(4, "little") for x in got) == want


def test_stream_is_hchacha_subkey_then_counter_blocks():
    """cuda/src/ChaCha.cu:109-166: key = HChaCha20(seed, domain || stream2); counter = stream0 (64-bit, carries); nonce = stream1."""
    seed = [0x0123456789ABCDEF, 0xFEDCBA9876543210, 0x1111111122222222, 0x3333333344444444]
    s0, s1, s2, dom = 0xFFFFFFFF, 0xAABBCCDD00112233, 0x5566778899AABBCC, 0x6F70656E66686531
    words = orc.stream_words(seed, s0, s1, s2, dom, 24)
    base_key = [w for x in seed for w in (x & 0xFFFFFFFF, x >> 32)]
    sub = orc.hchacha20(base_key, [dom & 0xFFFFFFFF, dom >> 32, s2 & 0xFFFFFFFF, s2 >> 32])
    for blk in range(3):
        ctr = s0 + blk  # crosses 2^32: the carry goes into word 13
        b = orc.chacha20_block(sub, [ctr & 0xFFFFFFFF, ctr >> 32, s1 & 0xFFFFFFFF, s1 >> 32])
        want = [int(b[2 * i]) | (int(b[2 * i + 1]) << 32) for i in range(8)]
        assert [int(x) for x in words[8 * blk:8 * blk + 8]] == want


def test_karney_moments():
    for mean, sd in ((0.0, SIGMA), (0.37, 1.5), (-12.25, 300.0), (5.5, 0.8)):
        z = orc.karney_samples([9, 8, 7, 6], mean, sd, 120000).astype(np.float64)
        assert abs(z.mean() - mean) < 5 * sd / np.sqrt(z.size) + 1e-9
        # a discrete Gaussian's variance equals sd^2 up to e^{-2 pi^2 sd^2}-sized terms for sd >~ 0.8
        assert abs(z.std() / sd - 1) < 0.02
    assert orc.karney_samples([1, 0, 0, 0], 3.4, 0.0, 4).tolist() == [3] * 4  # degenerate: llround(mean)


# ---------------------------------------------------------------- oracle: distribution properties
@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_oracle_distributions(n, depth, bits, bb):
    plan = plan_for(n, depth, bits)
    seed = [3, 1, 4, 1]
    u = plan.sample_distribution(2, 3, DIST_UNIFORM, 0.0, seed)
    assert all((u[:, :, l, :] < plan.moduli[l]).all() for l in range(plan.L))
    assert len(np.unique(u[0, 0, 0])) > n // 2  # independent streams per coefficient
    g = centred(plan.sample_distribution(2, 3, DIST_GAUSS, SIGMA, seed), plan.moduli)
    assert (g == g[:, :, :1, :]).all(), "the same integer in every limb"
    assert np.abs(g).max() < 12 * SIGMA
    b = plan.sample_distribution(2, 3, DIST_BIT, 0.0, seed)
    assert set(np.unique(b).tolist()) <= {0, 1}
    t = centred(plan.sample_distribution(2, 3, DIST_TERNARY, 0.0, seed), plan.moduli)
    assert set(np.unique(t).tolist()) <= {-1, 0, 1} and (t == t[:, :, :1, :]).all()
    # a column range of the wider matrix equals those columns of the whole (src/sampler/gpu.rs:197-250)
    for dist, sg in ((DIST_UNIFORM, 0.0), (DIST_GAUSS, SIGMA), (DIST_BIT, 0.0), (DIST_TERNARY, 0.0)):
        whole = plan.sample_distribution(2, 5, dist, sg, seed)
        part = plan.sample_distribution(2, 2, dist, sg, seed, full_cols=5, col_offset=2)
        assert np.array_equal(part, whole[:, 2:4])


@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_oracle_gauss_samp_relation(n, depth, bits, bb):
    """G z = v (mod q_t) per tower and coefficient, and a Gaussian-sized z (gpu_dcrt_poly.rs:2383-2540)."""
    plan = plan_for(n, depth, bits)
    src = plan.uniform(2, 2, seed=77)
    c = ((1 << bb) + 1.0) * SIGMA
    z = plan.gauss_samp_gq(src, bb, c, [5, 6, 7, 8])
    assert np.array_equal(gadget_recompose(plan, z, bb, 2), src.astype(object))
    zc = centred(z, plan.moduli)
    assert (zc == zc[:, :, :1, :]).all()
    assert np.abs(zc).max() < 12 * c
    assert not np.array_equal(z, plan.gauss_samp_gq(src, bb, c, [5, 6, 7, 9]))


def _p1_inputs(plan, d, bb, cols, seed0):
    k = plan.ell(bb)
    r = centred(plan.sample_distribution(d, d * k, DIST_GAUSS, SIGMA, [seed0, 0, 0, 1]), plan.moduli)[:, :, 0, :]
    e = centred(plan.sample_distribution(d, d * k, DIST_GAUSS, SIGMA, [seed0, 0, 0, 2]), plan.moduli)[:, :, 0, :]

    def gram(x, y):  # x y^T over Z[X]/(X^n+1), reduced into every limb
        n = plan.n
        out = np.zeros((d, d, plan.L, n), dtype=np.uint64)
        for i in range(d):
            for j in range(d):
                acc = np.zeros(n, dtype=object)
                for t in range(x.shape[1]):
                    full = np.convolve(x[i, t].astype(object), y[j, t].astype(object))
                    acc[: n] += full[: n]
                    acc[: n - 1] -= full[n:]
                for l, q in enumerate(plan.moduli):
                    out[i, j, l] = np.array([int(v) % q for v in acc], dtype=np.uint64)
        return out

    base = 1 << bb
    c = (base + 1.0) * SIGMA
    s = 1.8 * (base + 1.0) * SIGMA * SIGMA * (np.sqrt(d * plan.n * k) + np.sqrt(2 * plan.n) + 4.7)
    tp2 = plan.uniform(2 * d, cols, seed=seed0 + 5)
    # keep tp2 small and limb-consistent like the real [R;E] p2: centred Gaussian-sized integers
    small = centred(plan.sample_distribution(2 * d, cols, DIST_GAUSS, 40.0 * SIGMA, [seed0, 0, 0, 3]), plan.moduli)[:, :, 0, :]
    for l, q in enumerate(plan.moduli):
        tp2[:, :, l, :] = np.mod(small, q).astype(np.uint64)
    return gram(r, r), gram(r, e), gram(e, e), tp2, c, s


def test_oracle_p1_statistics():
    """Conditional sampling reproduces the covariance s^2 I - c^2 [[A, B], [B^T, D]] per coefficient (MatrixTrapdoor.cu:95-198)."""
    plan = plan_for(16, 2, 51)
    bb, d = 12, 1
    a, b, dm, tp2, c, s = _p1_inputs(plan, d, bb, 1, 21)
    tp2[:] = 0  # zero mean
    draws = np.stack([centred(plan.sample_p1(a, b, dm, tp2, c, s, SIGMA, [k, 1, 2, 3]), plan.moduli)[:, 0, 0, :] for k in range(400)])
    # draws: (trials, 2d, n)
    ac, bc, dc = (centred(x, plan.moduli)[0, 0, 0, :].astype(np.float64) for x in (a, b, dm))
    var0 = s * s - c * c * ac
    var1 = s * s - c * c * dc
    cov01 = -c * c * bc
    z = draws.astype(np.float64)
    est_var0, est_var1 = (z[:, 0, :] ** 2).mean(), (z[:, 1, :] ** 2).mean()
    assert abs(est_var0 / var0.mean() - 1) < 0.08 and abs(est_var1 / var1.mean() - 1) < 0.08
    est_cov = (z[:, 0, :] * z[:, 1, :]).mean(axis=0)
    assert np.abs(est_cov - cov01).max() < 0.35 * np.sqrt(var0.max() * var1.max())


# ---------------------------------------------------------------- product code on the host (emulation harness) vs oracle
@pytest.fixture(scope="module")
def emu():
    if not os.path.exists(EMU):
        import __graft_entry__ as ge

        ge.build()
    lib = C.CDLL(EMU)
    lib.emu_stream_words.argtypes = [u64p] + [C.c_uint64] * 4 + [C.c_size_t, u64p]
    lib.emu_stream_words.restype = None
    lib.emu_sample_distribution.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, C.c_int, C.c_double, u64p] + [C.c_size_t] * 4 + [u64p]
    lib.emu_gauss_samp.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, u64p, C.c_size_t, C.c_size_t, C.c_uint32, C.c_uint32,
                                   C.c_double, u64p, u64p]
    lib.emu_sample_p1.argtypes = [C.c_int, C.c_int, C.c_uint32, u64p, u64p, u64p, u64p, u64p, C.c_uint32, C.c_size_t,
                                  C.c_double, C.c_double, C.c_double, u64p, u64p]
    return lib


def test_emu_stream_matches_oracle(emu):
    seed = orc.seed_words([11, 22, 33, 44])
    for s0, s1, s2, dom in ((1, 1, 0, 0x6F70656E66686532), (0xFFFFFFFE, 7, 9, 0x6761646765746731), (2**40, 2**63, 2**33 + 5, 0x7065727475726231)):
        out = np.zeros(40, dtype=np.uint64)
        emu.emu_stream_words(p(seed), s0, s1, s2, dom, 40, p(out))
        assert np.array_equal(out, orc.stream_words(seed, s0, s1, s2, dom, 40))


@pytest.mark.parametrize("n,depth,bits,bb", CASES)
@pytest.mark.parametrize("wide", [0, 1])
def test_emu_distributions_bit_exact(emu, n, depth, bits, bb, wide):
    plan = plan_for(n, depth, bits)
    mod, seed = np.array(plan.moduli, dtype=np.uint64), orc.seed_words([2, 7, 1, 8])
    for dist, sg in ((DIST_UNIFORM, 0.0), (DIST_GAUSS, SIGMA), (DIST_GAUSS, 2.5e7), (DIST_BIT, 0.0), (DIST_TERNARY, 0.0)):
        got = np.zeros((2, 2, plan.L, n), dtype=np.uint64)
        assert emu.emu_sample_distribution(wide, n.bit_length() - 1, plan.L, p(mod), dist, sg, p(seed), 2, 2, 5, 1, p(got)) == 0
        assert np.array_equal(got, plan.sample_distribution(2, 2, dist, sg, seed, full_cols=5, col_offset=1)), (dist, sg)


@pytest.mark.parametrize("n,depth,bits,bb", CASES)
def test_emu_gauss_samp_relation(emu, n, depth, bits, bb):
    plan = plan_for(n, depth, bits)
    mod, seed = np.array(plan.moduli, dtype=np.uint64), orc.seed_words([5, 6, 7, 8])
    src = plan.uniform(2, 2, seed=78)
    c = ((1 << bb) + 1.0) * SIGMA
    z = np.zeros((2 * plan.ell(bb), 2, plan.L, n), dtype=np.uint64)
    assert emu.emu_gauss_samp(0, n.bit_length() - 1, plan.L, p(mod), p(src), 2, 2, bb, plan.dpt(bb), c, p(seed), p(z)) == 0
    assert np.array_equal(gadget_recompose(plan, z, bb, 2), src.astype(object))
    zc = centred(z, plan.moduli)
    assert (zc == zc[:, :, :1, :]).all() and np.abs(zc).max() < 12 * c
    # same streams, same algorithm: the host build of the product code and the oracle agree except where libm's
    # log/cos or a fused multiply-add rounds the continuous perturbation differently (rare)
    agree = (z == plan.gauss_samp_gq(src, bb, c, seed)).mean()
    assert agree > 0.98, agree


def test_emu_p1_matches_oracle(emu):
    plan = plan_for(16, 2, 51)
    bb, d, cols = 12, 2, 3
    a, b, dm, tp2, c, s = _p1_inputs(plan, d, bb, cols, 31)
    mod, seed = np.array(plan.moduli, dtype=np.uint64), orc.seed_words([4, 3, 2, 1])
    got = np.zeros_like(tp2)
    assert emu.emu_sample_p1(0, 4, plan.L, p(mod), p(a), p(b), p(dm), p(tp2), d, cols, c, s, SIGMA, p(seed), p(got)) == 0
    want = plan.sample_p1(a, b, dm, tp2, c, s, SIGMA, seed)
    assert (got == want).mean() > 0.98
    zc = centred(got, plan.moduli)
    assert (zc == zc[:, :, :1, :]).all() and np.abs(zc).max() < 12 * s


# ---------------------------------------------------------------- lane-balanced drivers (K tasks per thread): same samples
def _batched_emu(emu):
    emu.emu_sample_gauss_batched.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint32, u64p, C.c_double, u64p] + [C.c_size_t] * 4 + [u64p]
    emu.emu_gauss_samp_words.argtypes = [C.c_int, C.c_uint32, u64p, u64p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_double, u64p,
                                         C.POINTER(C.c_uint32)]
    return emu


@pytest.mark.parametrize("n,depth,bits,bb", [(256, 3, 24, 12), (256, 2, 24, 5), (256, 2, 51, 17), (256, 2, 59, 8), (512, 2, 28, 14)])
def test_emu_batched_samplers_draw_the_same_samples(emu, n, depth, bits, bb):
    """gauss_integers_batch (a thread runs K tasks through one word-per-trip loop and resumes each stream on a key-stream block
    boundary) against the one-task form, through the kernel's own thread body."""
    emu = _batched_emu(emu)
    plan = plan_for(n, depth, bits)
    mod, seed = np.array(plan.moduli, dtype=np.uint64), orc.seed_words([5, 6, 7, bb])
    logn = n.bit_length() - 1
    for sg in (SIGMA, 2.5e7, 0.3):
        want = np.zeros((2, 2, plan.L, n), dtype=np.uint64)
        assert emu.emu_sample_distribution(0, logn, plan.L, p(mod), DIST_GAUSS, sg, p(seed), 2, 2, 5, 1, p(want)) == 0
        for k in (2, 4, 8):
            got = np.zeros_like(want)
            assert emu.emu_sample_gauss_batched(k, 0, logn, plan.L, p(mod), sg, p(seed), 2, 2, 5, 1, p(got)) == 0
            assert np.array_equal(want, got), (sg, k)


def test_warp_model_of_the_g_lattice_sampler(emu):
    """The numbers DESIGN.md quotes for the sampler's lane utilisation: stream words per (poly, tower, coeff) task of the
    C5-shaped G-lattice sampler (two digits per tower), a warp of one-task lanes runs max-over-lanes trips, a warp of
    K-task lanes runs the longest sum (each task starts on a multiple of 8 trips)."""
    emu = _batched_emu(emu)
    n, depth, bits, bb = 4096, 3, 24, 12
    plan = plan_for(n, depth, bits)
    mod, seed = np.array(plan.moduli, dtype=np.uint64), orc.seed_words([5, 6, 7, 8])
    src = plan.uniform(1, 2, seed=78)
    w = np.zeros(src.size, dtype=np.uint32)
    c = ((1 << bb) + 1.0) * SIGMA
    assert emu.emu_gauss_samp_words(12, plan.L, p(mod), p(src), 2, bb, plan.dpt(bb), c, p(seed), w.ctypes.data_as(C.POINTER(C.c_uint32))) == 0
    assert 35 < w.mean() < 50 and w.min() >= 4 + 2 * 3  # two normals (4 words), two integer draws

    def busy_lanes(k):
        warps = w.size // (32 * k)
        per = w[: warps * 32 * k].reshape(warps, k, 32).astype(np.int64)
        t = np.zeros((warps, 32), dtype=np.int64)
        for i in range(k):
            t = (t + 7) // 8 * 8 + per[:, i, :]
        return 32.0 * per.sum() / (32.0 * t.max(1).sum())

    lanes = {k: busy_lanes(k) for k in (1, 4, 8)}
    assert 11 < lanes[1] < 14 and 17 < lanes[4] < 20 and 19.5 < lanes[8] < 23, lanes


# ---------------------------------------------------------------- golden vectors from the reference's CUDA library
def _golden():
    if not os.path.exists(GOLDEN):
        pytest.skip("tests/golden/ref_cuda_sampler_vectors.npz not generated yet")
    return np.load(GOLDEN)


def golden_cases():
    from tests.golden.make_golden_samplers import CASES as GC

    return GC


@pytest.mark.parametrize("case", range(4))
def test_oracle_matches_reference_cuda_samplers(case):
    """Integer-exact distributions: the oracle reproduces the reference library's matrices bit for bit."""
    g = _golden()
    n, depth, bits, bb = golden_cases()[case]
    plan = plan_for(n, depth, bits)
    key = f"n{n}_d{depth}_b{bits}_bb{bb}"
    seed = [int(x) for x in g[f"{key}/seed"]]
    for name, dist, sg in (("uniform", DIST_UNIFORM, 0.0), ("gauss", DIST_GAUSS, SIGMA), ("gauss_big", DIST_GAUSS, 3.1e7),
                           ("bit", DIST_BIT, 0.0), ("ternary", DIST_TERNARY, 0.0)):
        assert np.array_equal(plan.ntt(plan.sample_distribution(2, 3, dist, sg, seed)), g[f"{key}/{name}"]), name
        assert np.array_equal(plan.ntt(plan.sample_distribution(2, 3, dist, sg, seed, full_cols=5, col_offset=1)), g[f"{key}/{name}_cols"])
    # the reference's own G-lattice output satisfies the relation and agrees with the oracle almost everywhere
    z_ref = plan.intt(g[f"{key}/gauss_samp"])
    src = g[f"{key}/gauss_src"]
    assert np.array_equal(gadget_recompose(plan, z_ref, bb, src.shape[0]), src.astype(object))
    c = ((1 << bb) + 1.0) * SIGMA
    assert (z_ref == plan.gauss_samp_gq(src, bb, c, seed)).mean() > 0.98


# ---------------------------------------------------------------- GPU: the product through the C ABI
def _gpu_params(n, depth, bits, bb):
    from mxx_b200 import GpuDCRTPolyParams

    moduli = orc.prime_chain(n, bits, depth)
    return GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", CASES + [(4096, 3, 24, 12), (1024, 2, 51, 17)])
def test_gpu_distributions_bit_exact_vs_oracle(n, depth, bits, bb):
    from mxx_b200 import GpuDCRTPolyMatrix as M

    params, plan = _gpu_params(n, depth, bits, bb)
    seed = [0xDEADBEEF, 0x1234, 0x5678, 0x9ABC]
    for dist, sg in ((DIST_UNIFORM, 0.0), (DIST_GAUSS, SIGMA), (DIST_GAUSS, 2.5e7), (DIST_BIT, 0.0), (DIST_TERNARY, 0.0)):
        got = M.sample_distribution(params, 2, 3, dist, sg, seed)
        assert got.is_ntt
        assert np.array_equal(got.to_rns(), plan.ntt(plan.sample_distribution(2, 3, dist, sg, seed))), (dist, sg)
        part = M.sample_distribution_columns(params, 2, 7, 3, 2, dist, sg, seed)
        assert np.array_equal(part.to_rns(), plan.ntt(plan.sample_distribution(2, 2, dist, sg, seed, full_cols=7, col_offset=3)))


@pytest.mark.gpu
def test_gpu_sampler_error_paths():
    from mxx_b200 import GpuDCRTPolyMatrix as M, GpuError

    params, _ = _gpu_params(64, 2, 24, 12)
    with pytest.raises(GpuError, match="dist_type"):
        M.sample_distribution(params, 1, 1, 7, 1.0, 1)
    with pytest.raises(GpuError, match="sigma"):
        M.sample_distribution(params, 1, 1, DIST_GAUSS, 0.0, 1)
    out = M.new_empty(params, 1, 3)
    st = out._L.gpu_matrix_sample_distribution_columns(out.raw, DIST_BIT, 0.0, __import__("mxx_b200.matrix", fromlist=["as_seed"]).as_seed(1), 4, 2)
    assert st != 0 and b"column range" in out._L.gpu_last_error()
    src = M.zero(params, 1, 1)
    with pytest.raises(GpuError, match="c must be positive"):
        src.gauss_samp_gq_arb_base(0.0, SIGMA, 1)
    assert M.sample_distribution(params, 0, 3, DIST_BIT, 0.0, 1).size() == (0, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb", CASES + [(2048, 3, 24, 12)])
def test_gpu_gauss_samp_relation(n, depth, bits, bb):
    """test_gpu_matrix_gauss_samp_gq_arb_base_relation (gpu_dcrt_poly.rs:2383-2540): G * sampled == source, 16 seeds."""
    from mxx_b200 import GpuDCRTPolyMatrix as M

    params, plan = _gpu_params(n, depth, bits, bb)
    c = ((1 << bb) + 1.0) * SIGMA
    src_h = plan.uniform(3, 2, seed=123)
    src_eval = M.from_rns(params, plan.ntt(src_h), True)
    gadget = M.gadget_matrix(params, 3)
    for off in range(16 if n <= 128 else 3):
        z = src_eval.clone().gauss_samp_gq_arb_base(c, SIGMA, [0x123456789ABCDEF0 + off, 0, 0, 0])
        assert z.is_ntt and z.size() == (3 * params.modulus_digits(), 2)
        assert gadget * z == src_eval
        zc = centred(plan.intt(z.to_rns()), plan.moduli)
        assert (zc == zc[:, :, :1, :]).all() and np.abs(zc).max() < 12 * c
    # same seed -> same matrix; agrees with the oracle except for libm / FMA rounding of the continuous perturbation
    z1 = src_eval.clone().gauss_samp_gq_arb_base(c, SIGMA, [1, 2, 3, 4]).to_rns()
    z2 = src_eval.clone().gauss_samp_gq_arb_base(c, SIGMA, [1, 2, 3, 4]).to_rns()
    assert np.array_equal(z1, z2)
    assert (plan.intt(z1) == plan.gauss_samp_gq(src_h, bb, c, [1, 2, 3, 4])).mean() > 0.98


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(4))
def test_gpu_matches_reference_cuda_samplers(case):
    """Every sampler, sample for sample, against matrices the reference's CUDA library produced on a B200."""
    from mxx_b200 import GpuDCRTPolyMatrix as M
    from tests.golden.make_golden_samplers import run_case

    g = _golden()
    n, depth, bits, bb = golden_cases()[case]
    key = f"n{n}_d{depth}_b{bits}_bb{bb}"
    mine = run_case(None, n, depth, bits, bb)
    assert mine, "no outputs"
    for k, v in mine.items():
        assert np.array_equal(v, g[f"{key}/{k}"]), k
    assert M is not None


BATCH_KNOB_SCRIPT = """
import sys
import numpy as np
sys.path.insert(0, {root!r})
from mxx_b200 import GpuDCRTPolyMatrix as M, GpuDCRTPolyParams
from mxx_b200.matrix import DIST_GAUSS
from oracle import oracle as orc
out = {{}}
for n, depth, bits, bb in [(1024, 3, 24, 12), (512, 2, 51, 17), (256, 2, 28, 7)]:
    moduli = orc.prime_chain(n, bits, depth)
    params, plan = GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)
    seed = [n, bits, 3, 4]
    key = "n%d_b%d" % (n, bits)
    out[key + "_gauss"] = M.sample_distribution(params, 2, 3, DIST_GAUSS, 4.578, seed).to_rns()
    out[key + "_gauss_big"] = M.sample_distribution_columns(params, 2, 5, 1, 3, DIST_GAUSS, 3.1e7, seed).to_rns()
    src = plan.uniform(2, 2, seed=9)
    out[key + "_gq"] = M.from_rns(params, src, False).gauss_samp_gq_arb_base(((1 << bb) + 1.0) * 4.578, 4.578, seed).to_rns()
np.savez({path!r}, **out)
print("sampler knob ok")
"""


@pytest.mark.gpu
def test_gpu_sampler_batch_knob_keeps_samples(tmp_path):
    """MXX_SAMPLER_BATCH = 1 (one task per thread) / 2 / 4 / 8 tasks per lane: the same matrices, word for word."""
    import subprocess

    got = {}
    for k in ("1", "2", "4", "8"):
        path = str(tmp_path / f"k{k}.npz")
        run = subprocess.run([sys.executable, "-c", BATCH_KNOB_SCRIPT.format(root=ROOT, path=path)], env=dict(os.environ, MXX_SAMPLER_BATCH=k),
                             capture_output=True, text=True, timeout=900)
        assert run.returncode == 0 and "sampler knob ok" in run.stdout, run.stderr[-2000:]
        got[k] = dict(np.load(path))
    for k in ("2", "4", "8"):
        for name, v in got["1"].items():
            assert np.array_equal(v, got[k][name]), (k, name)


@pytest.mark.gpu
def test_gpu_vs_reference_library_side_by_side():
    """Fresh seeds, both libraries in this process (skipped when oracle/_ref was not built)."""
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libgpupoly_ref.so")
    if not os.path.exists(ref_path):
        pytest.skip("oracle/_ref/libgpupoly_ref.so not built")
    from mxx_b200 import ffi
    from tests.golden.make_golden_samplers import run_case

    ref = ffi.bind(ref_path, extensions=False)
    for case in ((256, 2, 28, 7), (128, 3, 51, 13)):
        mine, theirs = run_case(None, *case, seed0=0xABCDEF), run_case(ref, *case, seed0=0xABCDEF)
        for k in mine:
            assert np.array_equal(mine[k], theirs[k]), (case, k)


@pytest.mark.gpu
@pytest.mark.parametrize("n,depth,bits,bb,d,cols", [(128, 2, 16, 8, 1, 1), (128, 2, 16, 8, 2, 3), (64, 3, 24, 12, 1, 2), (256, 2, 51, 17, 2, 2),
                                                    (64, 2, 24, 12, 5, 2)])
def test_gpu_trapdoor_preimage(n, depth, bits, bb, d, cols):
    """src/sampler/trapdoor/gpu.rs tests: A * preimage(A, T, u) == u, and the preimage is short."""
    from mxx_b200 import GpuDCRTPolyMatrix as M
    from mxx_b200.trapdoor import GpuDCRTPolyTrapdoorSampler, preimage_smoothing_parameter

    params, plan = _gpu_params(n, depth, bits, bb)
    sampler = GpuDCRTPolyTrapdoorSampler(params, SIGMA)
    td, a = sampler.trapdoor(params, d, seed=[1, 2, 3, d])
    k = params.modulus_digits()
    assert a.size() == (d, (k + 2) * d)
    target = M.sample_distribution(params, d, cols, DIST_UNIFORM, 0.0, [9, 9, 9, cols])
    x = sampler.preimage(params, td, a, target, seed=[4, 5, 6, 7])
    assert x.size() == ((k + 2) * d, cols)
    assert a * x == target
    s = preimage_smoothing_parameter(1 << bb, SIGMA, d, n, k)
    if s * 12 < min(plan.moduli) / 2:
        xc = centred(plan.intt(x.to_rns()), plan.moduli)
        assert (xc == xc[:, :, :1, :]).all(), "the preimage is one short integer vector, identical in every limb"
        assert np.abs(xc).max() < 12 * s * np.sqrt(d * k * n) / np.sqrt(n)  # loose: |R z| <= sqrt(dk) * 12 sigma * 12 c
    # a second seed gives a different preimage of the same target
    x2 = sampler.preimage(params, td, a, target, seed=[4, 5, 6, 8])
    assert a * x2 == target and not (x2 == x)


@pytest.mark.gpu
def test_gpu_preimage_extend_and_trapdoor_bytes():
    from mxx_b200 import GpuDCRTPolyMatrix as M
    from mxx_b200.trapdoor import GpuDCRTPolyTrapdoorSampler, GpuDCRTTrapdoor

    params, _ = _gpu_params(128, 2, 24, 12)
    sampler = GpuDCRTPolyTrapdoorSampler(params, SIGMA)
    td, a = sampler.trapdoor(params, 2, seed=[7, 7, 7, 7])
    ext = M.sample_distribution(params, 2, 3, DIST_UNIFORM, 0.0, [8, 0, 0, 0])
    target = M.sample_distribution(params, 2, 2, DIST_UNIFORM, 0.0, [9, 0, 0, 0])
    x = sampler.preimage_extend(params, td, a, ext, target, seed=[1, 1, 1, 1])
    assert a.concat_columns([ext]) * x == target
    td2 = GpuDCRTTrapdoor.from_compact_bytes(params, td.to_compact_bytes(), SIGMA)
    assert td2 is not None and td2.r == td.r and td2.e == td.e
    assert td2.a_mat_coeff == td.a_mat_coeff and td2.b_mat_coeff == td.b_mat_coeff and td2.d_mat_coeff == td.d_mat_coeff
    y = sampler.preimage(params, td2, a, target, seed=[2, 2, 2, 2])
    assert a * y == target
