"""GPU parity tests: the CUDA path, called through the C ABI (libgpupoly.so via
mxx_b200.ffi), against the CPU oracle on the same seeded inputs — bit-exact residues per
limb and per coefficient — plus the reference's own relational tests
(src/matrix/gpu_dcrt_poly.rs:2036-2700, src/matrix/dcrt_poly.rs:515-799).
"""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mx():
    import mxx_b200

    return mxx_b200


_cache = {}


def setup(mx, n, depth, bits, bb):
    key = (n, depth, bits, bb)
    if key not in _cache:
        moduli = mx.standard_moduli(n, depth, bits)
        assert moduli == orc.prime_chain(n, bits, depth)
        _cache[key] = (mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli))
    return _cache[key]


M = None  # set per test from the fixture


def G(mx, params, host, is_ntt=True):
    return mx.GpuDCRTPolyMatrix.from_rns(params, host, is_ntt)


# (n, depth, bits, base_bits): u32 path (24/28/29-bit), u64 path (51/61-bit), tiny rings, multi-pass n
PARAMS = [
    (4, 2, 17, 1),
    (4, 1, 52, 17),
    (128, 2, 17, 3),
    (1024, 3, 24, 12),
    (4096, 2, 29, 10),
    (2048, 2, 51, 17),
    (1024, 2, 61, 20),
    (16384, 2, 24, 12),
    (16384, 2, 51, 17),
    (32768, 1, 28, 14),
    (32768, 1, 51, 17),
    (65536, 1, 28, 14),
    (65536, 1, 51, 26),
    (16384, 2, 55, 19),   # u64 integer kernels without conditional subtractions (moduli in [2^51, 2^58): not on the FP64 pipe)
    (32768, 1, 55, 19),
]


@pytest.mark.parametrize("n,depth,bits,bb", PARAMS)
def test_ntt_intt_match_oracle(mx, n, depth, bits, bb):
    params, plan = setup(mx, n, depth, bits, bb)
    a = plan.uniform(2, 2, seed=n + bits)
    g = G(mx, params, a, is_ntt=False)
    assert np.array_equal(g.to_rns(), a)  # load/store round trip, coefficient domain
    g.ntt_in_place()
    ev = g.to_rns()
    assert np.array_equal(ev, plan.ntt(a))  # natural-order boundary layout X[k] = a(psi^(2k+1))
    g.intt_all_in_place()
    assert np.array_equal(g.to_rns(), a)
    # EVAL-tagged load + INTT
    h = G(mx, params, ev, is_ntt=True)
    h.intt_all_in_place()
    assert np.array_equal(h.to_rns(), a)


@pytest.mark.parametrize("n,depth,bits,bb", PARAMS[:9])
@pytest.mark.parametrize("shape", [(1, 5, 7), (3, 9, 5), (5, 4, 6), (2, 33, 1), (1, 1, 1)])
def test_mul_matches_oracle(mx, n, depth, bits, bb, shape):
    if n >= 16384 and shape != (1, 5, 7):
        pytest.skip("large ring: one shape is enough")
    params, plan = setup(mx, n, depth, bits, bb)
    r, k, c = shape
    A, B = plan.uniform(r, k, seed=1), plan.uniform(k, c, seed=2)
    got = (G(mx, params, A) * G(mx, params, B)).to_rns()
    assert np.array_equal(got, plan.mul(A, B))


def test_mul_is_negacyclic_convolution(mx):
    """coefficient-domain definition: INTT(NTT(a) . NTT(b)) == a*b mod (x^n+1, q) by schoolbook"""
    n, depth, bits, bb = 256, 2, 28, 14
    params, plan = setup(mx, n, depth, bits, bb)
    a, b = plan.uniform(1, 1, seed=3), plan.uniform(1, 1, seed=4)
    ga, gb = G(mx, params, a, False), G(mx, params, b, False)
    prod = ga * gb  # converts copies to the evaluation domain (gpu.rs:1122-1141)
    prod.intt_all_in_place()
    got = prod.to_rns()
    for l, q in enumerate(plan.moduli):
        want = orc.negacyclic_schoolbook(a[0, 0, l], b[0, 0, l], q)
        assert np.array_equal(got[0, 0, l], want)


@pytest.mark.parametrize("n,depth,bits,bb", PARAMS[:7])
def test_add_sub_scalar_neg(mx, n, depth, bits, bb):
    params, plan = setup(mx, n, depth, bits, bb)
    A, B = plan.uniform(2, 3, seed=5), plan.uniform(2, 3, seed=6)
    s = plan.uniform(1, 1, seed=7)
    gA, gB, gs = G(mx, params, A), G(mx, params, B), G(mx, params, s)
    assert np.array_equal((gA + gB).to_rns(), plan.add(A, B))
    assert np.array_equal((gA - gB).to_rns(), plan.sub(A, B))
    assert np.array_equal((gA * gs).to_rns(), plan.mul_scalar(A, s))
    assert np.array_equal((-gA).to_rns(), plan.sub(np.zeros_like(A), A))
    assert (gA - gB) + gB == gA  # src/poly/dcrt/poly.rs:631-720 ring identities
    acc = gA.clone()
    acc.add_in_place(gB)
    assert acc == gA + gB
    assert not (gA == gB)


@pytest.mark.parametrize("n,depth,bits,bb", PARAMS[:9] + [PARAMS[10]])
@pytest.mark.parametrize("small", [False, True])
def test_decompose_matches_oracle_and_gadget_relation(mx, n, depth, bits, bb, small):
    params, plan = setup(mx, n, depth, bits, bb)
    rows, cols = (2, 3) if n < 16384 else (1, 2)
    Mh = plan.uniform(rows, cols, seed=8)  # EVAL-tagged input, as BggPublicKey holds it
    gM = G(mx, params, Mh)
    dec = gM.small_decompose() if small else gM.decompose()
    want = plan.ntt(plan.decompose(plan.intt(Mh), bb, small))
    assert np.array_equal(dec.to_rns(), want)
    # coefficient-domain input gives the same digits
    gC = G(mx, params, plan.intt(Mh), is_ntt=False)
    assert (gC.small_decompose() if small else gC.decompose()) == dec
    if not small:
        # G * G^-1(M) == M  (gpu_dcrt_poly.rs:2077-2163, dcrt_poly.rs:515-600, 683-728)
        gad = mx.GpuDCRTPolyMatrix.gadget_matrix(params, rows)
        assert gad.size() == (rows, rows * params.modulus_digits())
        assert gad * dec == gM
        assert np.array_equal(gad.to_rns(), plan.ntt(plan.gadget(rows, bb)))
    else:
        sg = mx.GpuDCRTPolyMatrix.small_gadget_matrix(params, rows)
        assert np.array_equal(sg.to_rns(), plan.ntt(plan.gadget(rows, bb, small=True)))


def test_decompose_chunk_matches_full(mx):
    """gpu_dcrt_poly.rs:2167-2221"""
    params, plan = setup(mx, 128, 2, 17, 3)
    gM = G(mx, params, plan.uniform(2, 4, seed=9))
    full = gM.decompose()
    for c0, c1 in [(0, 1), (1, 3), (3, 4)]:
        assert gM.slice_columns(c0, c1).decompose() == full.slice_columns(c0, c1)
        assert gM.slice_columns(c0, c1).small_decompose() == gM.small_decompose().slice_columns(c0, c1)


MD = [
    # (n, depth, bits, bb, r, d, c)
    (4, 2, 17, 1, 1, 1, 2),
    (128, 2, 17, 3, 2, 2, 3),
    (1024, 3, 24, 12, 1, 1, 4),
    (1024, 3, 24, 12, 4, 2, 3),
    (2048, 2, 51, 17, 2, 2, 2),
    (16384, 3, 24, 12, 1, 1, 3),
    (16384, 2, 51, 17, 1, 1, 2),
    (32768, 1, 28, 14, 1, 1, 2),
    (32768, 1, 51, 17, 1, 1, 2),   # u64 ring above one block: digits extracted inside the first global pass
    (65536, 1, 28, 14, 1, 1, 1),
]


@pytest.mark.parametrize("n,depth,bits,bb,r,d,c", MD)
@pytest.mark.parametrize("small", [False, True])
def test_mul_decompose_matches_oracle(mx, n, depth, bits, bb, r, d, c, small):
    params, plan = setup(mx, n, depth, bits, bb)
    ell = plan.ell(bb, small)
    S, other = plan.uniform(r, d * ell, seed=10), plan.uniform(d, c, seed=11)
    want = plan.mul_decompose(S, other, bb, small)
    gS, gO = G(mx, params, S), G(mx, params, other)
    fused = gS.mul_decompose(gO, fused=True, small=small)
    seq = gS.mul_decompose(gO, fused=False, small=small)  # the reference's FFI call sequence
    assert np.array_equal(fused.to_rns(), want)
    assert np.array_equal(seq.to_rns(), want)
    assert gS * (gO.small_decompose() if small else gO.decompose()) == fused
    assert np.array_equal(gO.to_rns(), other)  # inputs untouched


KNOB_SCRIPT = """
import sys
import numpy as np
sys.path.insert(0, {root!r})
import mxx_b200 as mx
from oracle import oracle as orc
for n, depth, bits, bb in [(16384, 2, 51, 17), (32768, 1, 51, 17)]:
    moduli = mx.standard_moduli(n, depth, bits)
    params, plan = mx.GpuDCRTPolyParams(n, moduli, bb), orc.Plan(n, moduli)
    a = plan.uniform(1, 2, seed=3)
    g = mx.GpuDCRTPolyMatrix.from_rns(params, a, False)
    g.ntt_in_place()
    assert np.array_equal(g.to_rns(), plan.ntt(a))
    g.intt_all_in_place()
    assert np.array_equal(g.to_rns(), a)
    ell = plan.ell(bb, False)
    S, other = plan.uniform(2, ell, seed=10), plan.uniform(1, 2, seed=11)
    got = mx.GpuDCRTPolyMatrix.from_rns(params, S, True).mul_decompose(mx.GpuDCRTPolyMatrix.from_rns(params, other, True), fused=True)
    assert np.array_equal(got.to_rns(), plan.mul_decompose(S, other, bb, False))
print("knobs ok")
"""


@pytest.mark.parametrize("knobs", [{"MXX_NTT_F64": "0"}, {"MXX_MATMUL_F64": "0"}, {"MXX_DECOMP_FUSE_TOP": "0"},
                                   {"MXX_NTT_F64": "0", "MXX_MATMUL_F64": "0", "MXX_DECOMP_FUSE_TOP": "0"}])
def test_kernel_selection_knobs_keep_results(mx, knobs):
    """51-bit contexts default to the FP64-pipe NTT / product and the fused top pass; with each knob off the integer
    kernels must give the same (oracle) results.  The knobs are read at context creation, hence a fresh process."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, **knobs)
    out = subprocess.run([sys.executable, "-c", KNOB_SCRIPT.format(root=root)], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "knobs ok" in out.stdout, out.stderr[-2000:]


def test_mul_decompose_chunked_workspace(mx, monkeypatch):
    """force several column chunks through the fused path"""
    params, plan = setup(mx, 1024, 3, 24, 12)
    ell = plan.ell(12)
    S, other = plan.uniform(2, ell, seed=12), plan.uniform(1, 5, seed=13)
    monkeypatch.setenv("MXX_MUL_DECOMPOSE_SCRATCH_BYTES", str(2 * ell * 3 * 1024 * 4))  # one column per chunk: five chunks over two lanes
    got = G(mx, params, S).mul_decompose(G(mx, params, other), fused=True)
    assert np.array_equal(got.to_rns(), plan.mul_decompose(S, other, 12))


def test_mul_decompose_small_relation(mx):
    """(A * G_small).mul_decompose_small(B) == A * B when B has small coefficients
    (gpu_dcrt_poly.rs:2336-2379, dcrt_poly.rs:754-799)"""
    n, depth, bits, bb = 128, 2, 17, 3
    params, plan = setup(mx, n, depth, bits, bb)
    A = plan.uniform(2, 2, seed=14)
    rng = np.random.default_rng(15)
    bound = 1 << bits  # coefficients below 2^crt_bits-1 are their own limb-0 digits' recomposition
    coeff = rng.integers(0, min(bound, min(plan.moduli)), size=(2, 3, 1, n), dtype=np.uint64)
    Bc = np.ascontiguousarray(np.broadcast_to(coeff, (2, 3, depth, n)))
    gA = G(mx, params, A)
    gB = G(mx, params, Bc, is_ntt=False)
    sg = mx.GpuDCRTPolyMatrix.small_gadget_matrix(params, 2)
    lhs = (gA * sg).mul_decompose_small(gB)
    assert lhs == gA * gB


def test_slice_concat_entry_transpose(mx):
    """gpu_dcrt_poly.rs:2546-2632"""
    params, plan = setup(mx, 128, 2, 17, 3)
    A, B = plan.uniform(2, 3, seed=16), plan.uniform(2, 2, seed=17)
    gA, gB = G(mx, params, A), G(mx, params, B)
    assert np.array_equal(gA.concat_columns([gB]).to_rns(), np.concatenate([A, B], axis=1))
    C2 = plan.uniform(1, 3, seed=18)
    assert np.array_equal(gA.concat_rows([G(mx, params, C2)]).to_rns(), np.concatenate([A, C2], axis=0))
    dg = gA.concat_diag([gB]).to_rns()
    assert np.array_equal(dg[:2, :3], A) and np.array_equal(dg[2:, 3:], B) and not dg[:2, 3:].any() and not dg[2:, :3].any()
    assert np.array_equal(gA.slice(1, 2, 1, 3).to_rns(), A[1:2, 1:3])
    assert np.array_equal(gA.entry(1, 2).to_rns(), A[1:2, 2:3])
    assert np.array_equal(gA.transpose().to_rns(), np.ascontiguousarray(A.transpose(1, 0, 2, 3)))
    acc = gA.clone()
    acc.add_block_from(gB, 0, 1, 0, 0, 2, 2)
    want = A.copy()
    want[:, 1:3] = plan.add(A[:, 1:3].copy(), B)
    assert np.array_equal(acc.to_rns(), want)
    ident = mx.GpuDCRTPolyMatrix.identity(params, 3)
    assert gA * ident == gA


def test_cross_domain_and_const_coeff(mx):
    params, plan = setup(mx, 128, 2, 17, 3)
    a = plan.uniform(1, 2, seed=19)
    g = G(mx, params, a, is_ntt=False)
    assert np.array_equal(g.store_const_coeff_words(), a[..., 0])
    e = g.clone()
    e.ntt_in_place()
    assert not (e == g)  # differing domains compare unequal without an error (MatrixArith.cu:2934-2937)
    with pytest.raises(mx.GpuError):
        e.store_const_coeff_words()  # requires the coefficient domain (MatrixSerde.cu:943)
    with pytest.raises(mx.GpuError):
        out = np.empty_like(a)
        e.store_rns(out, mx.GPU_POLY_FORMAT_COEFF)  # tag must equal the current domain (:765-768)


def test_level_below_context(mx):
    """matrices with fewer active limbs than the context (level < crt_depth-1)"""
    params, plan = setup(mx, 1024, 3, 24, 12)
    sub = orc.Plan(1024, plan.moduli[:2])
    A, B = sub.uniform(2, 2, seed=20), sub.uniform(2, 2, seed=21)
    gA, gB = G(mx, params, A), G(mx, params, B)
    assert gA.level == 1
    assert np.array_equal((gA * gB).to_rns(), sub.mul(A, B))
    c = G(mx, params, A, is_ntt=False)
    c.ntt_in_place()
    assert np.array_equal(c.to_rns(), sub.ntt(A))


def test_error_paths(mx):
    """validation rules of SURVEY §8b"""
    import ctypes as C

    from mxx_b200 import ffi

    L = ffi.lib()
    h = C.c_void_p()
    mod = (C.c_uint64 * 1)(97)
    ids = (C.c_int * 1)(0)
    assert L.gpu_context_create(5, 0, 0, mod, 1, ids, 1, C.byref(h)) != 0  # 97 != 1 mod 64
    assert b"2N" in L.gpu_last_error()
    assert L.gpu_context_create(3, 1, 0, mod, 1, ids, 1, C.byref(h)) != 0  # moduli_len != L+1
    params, plan = setup(mx, 128, 2, 17, 3)
    A = G(mx, params, plan.uniform(2, 3, seed=22))
    B = G(mx, params, plan.uniform(2, 3, seed=23))
    out = mx.GpuDCRTPolyMatrix(params, 2, 3)
    assert L.gpu_matrix_mul(out.raw, A.raw, B.raw) != 0  # inner dimension mismatch
    Bc = G(mx, params, plan.uniform(3, 3, seed=24), is_ntt=False)
    assert L.gpu_matrix_mul(out.raw, A.raw, Bc.raw) != 0  # requires EVAL (MatrixArith.cu:2863-2866)
    assert L.gpu_matrix_copy_block(out.raw, A.raw, 0, 0, 0, 2, 2, 2) != 0  # out of range
    seed = ffi.GpuRngSeed()
    assert L.gpu_matrix_sample_distribution(out.raw, 9, 0.0, seed) != 0  # invalid dist_type (MatrixSampling.cu:424-427)
    empty = mx.GpuDCRTPolyMatrix(params, 0, 3)
    assert empty.to_rns().shape == (0, 3, 2, 128)
    assert (mx.GpuDCRTPolyMatrix(params, 2, 0) * mx.GpuDCRTPolyMatrix(params, 0, 3)).to_rns().any() == False  # noqa: E712


@pytest.mark.parametrize("n,depth,bits,bb,r,d,c", [(16384, 15, 24, 12, 1, 1, 8), (16384, 4, 51, 17, 2, 2, 4)])
def test_full_size_properties(mx, n, depth, bits, bb, r, d, c):
    """BASELINE-size shapes checked through size-independent properties:
    G * G^-1(M) == M, linearity of mul_decompose in S, NTT round trip."""
    params, plan = setup(mx, n, depth, bits, bb)
    ell = plan.ell(bb)
    other = G(mx, params, plan.uniform(d, c, seed=30))
    S1, S2 = G(mx, params, plan.uniform(r, d * ell, seed=31)), G(mx, params, plan.uniform(r, d * ell, seed=32))
    assert mx.GpuDCRTPolyMatrix.gadget_matrix(params, d).mul_decompose(other) == other
    assert (S1 + S2).mul_decompose(other) == S1.mul_decompose(other) + S2.mul_decompose(other)
    assert S1.mul_decompose(other, fused=True) == S1.mul_decompose(other, fused=False)
    rt = other.clone()
    rt.intt_all_in_place()
    rt.ntt_in_place()
    assert rt == other
    # spot check of one output polynomial against the oracle
    one = S1.mul_decompose(other.slice_columns(0, 1)).to_rns()
    assert np.array_equal(one, plan.mul_decompose(S1.to_rns(), other.slice_columns(0, 1).to_rns(), bb))


@pytest.mark.parametrize("pinned", [False, True])
def test_large_load_chunked_pipeline(mx, pinned):
    """Loads above 128 MB on a u32 context go through the chunked two-lane pipeline (MatrixHostLoad.inl): host threads narrow
    chunks from the back while raw chunks are DMA'd from the front.  Same device contents as the single-DMA path, for both
    formats (EVAL applies the bit-reversal per chunk), with a ragged last chunk, from pageable and from pinned memory."""
    n, depth = 4096, 3
    params, plan = setup(mx, n, depth, 24, 12)
    rows, cols = 3, 487  # 1461 polys * 3 * 4096 = 17.95 M words: 8 full chunks + a tail
    host = plan.uniform(rows, cols, seed=4242)
    buf = None
    if pinned:
        buf = mx.PinnedBuffer(host.shape)
        buf.array[...] = host
        src = buf.array
    else:
        src = host
    for is_ntt in (False, True):
        m = mx.GpuDCRTPolyMatrix(params, rows, cols, depth - 1, is_ntt)
        m.load_rns(src, m._fmt())
        assert np.array_equal(m.to_rns(), host)
        # the chunks really are where the rest of the library expects them: transform + product on a slice
        if is_ntt:
            a = m.slice(0, 1, 0, 3)
            b = m.slice(0, 3, 484, 487)
            assert np.array_equal((a * b).to_rns(), plan.mul(host[0:1, 0:3], host[0:3, 484:487]))
        else:
            m.ntt_in_place()
            assert np.array_equal(m.slice(2, 3, 480, 487).to_rns(), plan.ntt(host[2:3, 480:487]))
    if buf is not None:
        buf.free()
